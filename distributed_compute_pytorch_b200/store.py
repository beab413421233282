"""Key-value store + ``env://`` rendezvous.

Covers what ``setup()`` obtains implicitly in the reference (ref:main.py:47-50): the process that is
rank 0 hosts a TCP store on ``MASTER_ADDR:MASTER_PORT`` and every rank connects to it.  The server
and client are native (csrc/host/store.cpp); this file is the thin Python face.
"""
from __future__ import annotations

import ctypes as C
import os
import pickle
import time
from typing import Any, Optional

from ._native import host_lib

DEFAULT_PORT = 12355  # ref:main.py:49
DEFAULT_TIMEOUT_S = 1800.0


class StoreTimeout(RuntimeError):
    pass


class TCPStore:
    """``TCPStore(host, port, world_size, is_master)`` with set/get/add/wait/check/delete_key."""

    def __init__(self, host_name: str, port: int, world_size: Optional[int] = None, is_master: bool = False,
                 timeout: float = DEFAULT_TIMEOUT_S, wait_for_workers: bool = False):
        self._lib = host_lib()
        self.host, self.world_size, self.is_master, self.timeout = host_name, world_size, is_master, float(timeout)
        self._server = None
        self.port = int(port)
        if is_master:
            out_port = C.c_int(0)
            self._server = self._lib.dcp_store_server_start(host_name.encode(), int(port), C.byref(out_port))
            if not self._server:
                raise RuntimeError(f"TCPStore: cannot listen on {host_name}:{port} (address in use?)")
            self.port = out_port.value
        self._client = self._lib.dcp_store_client_connect(host_name.encode(), self.port, self.timeout)
        if not self._client:
            self.close()
            raise StoreTimeout(f"TCPStore: cannot connect to {host_name}:{self.port}")
        if wait_for_workers and world_size:
            n = self.add("__workers__", 1)
            deadline = time.time() + self.timeout
            while n < world_size:
                if time.time() > deadline:
                    raise StoreTimeout("TCPStore: timed out waiting for workers")
                time.sleep(0.01)
                n = self.add("__workers__", 0)

    # -- raw bytes API ----------------------------------------------------------------------------
    def set(self, key: str, value) -> None:
        if isinstance(value, str):
            value = value.encode()
        value = bytes(value)
        buf = C.create_string_buffer(value, len(value))
        if self._lib.dcp_store_set(self._client, key.encode(), buf, len(value)) < 0:
            raise RuntimeError("TCPStore.set failed (connection lost)")

    def get(self, key: str, timeout: Optional[float] = None) -> bytes:
        t = self.timeout if timeout is None else float(timeout)
        n = self._lib.dcp_store_get(self._client, key.encode(), t)
        if n == -3:
            # a late reply would desynchronise this connection: reconnect
            self._lib.dcp_store_client_close(self._client)
            self._client = self._lib.dcp_store_client_connect(self.host.encode(), self.port, self.timeout)
            raise StoreTimeout(f"TCPStore.get({key!r}) timed out after {t}s")
        if n < 0:
            raise RuntimeError(f"TCPStore.get({key!r}) failed (connection lost)")
        buf = C.create_string_buffer(max(int(n), 1))
        self._lib.dcp_store_copy_last(self._client, buf, n)
        return buf.raw[:n]

    def add(self, key: str, amount: int) -> int:
        r = self._lib.dcp_store_add(self._client, key.encode(), int(amount))
        return int(r)

    def check(self, keys) -> bool:
        if isinstance(keys, str):
            keys = [keys]
        return all(self._lib.dcp_store_check(self._client, k.encode()) == 1 for k in keys)

    def wait(self, keys, timeout: Optional[float] = None) -> None:
        if isinstance(keys, str):
            keys = [keys]
        for k in keys:
            self.get(k, timeout)

    def delete_key(self, key: str) -> bool:
        return self._lib.dcp_store_delete(self._client, key.encode()) == 1

    def num_keys(self) -> int:
        return int(self._lib.dcp_store_num_keys(self._client))

    # -- pickled objects --------------------------------------------------------------------------
    def set_obj(self, key: str, obj: Any) -> None:
        self.set(key, pickle.dumps(obj))

    def get_obj(self, key: str, timeout: Optional[float] = None) -> Any:
        return pickle.loads(self.get(key, timeout))

    # -- host barrier over the store --------------------------------------------------------------
    def barrier(self, name: str, rank: int, world_size: int, timeout: Optional[float] = None) -> None:
        n = self.add(f"{name}/count", 1)
        if n == world_size:
            self.set(f"{name}/go", b"1")
        self.get(f"{name}/go", timeout)

    @property
    def client_handle(self):
        return self._client

    def close(self) -> None:
        if getattr(self, "_client", None):
            self._lib.dcp_store_client_close(self._client)
            self._client = None
        if getattr(self, "_server", None):
            self._lib.dcp_store_server_stop(self._server)
            self._server = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PrefixStore:
    """Namespaces keys so successive process groups on one store cannot collide."""

    def __init__(self, prefix: str, store: TCPStore):
        self.prefix, self.store = prefix, store

    def _k(self, key: str) -> str:
        return f"{self.prefix}/{key}"

    def set(self, key, value): self.store.set(self._k(key), value)
    def get(self, key, timeout=None): return self.store.get(self._k(key), timeout)
    def add(self, key, amount): return self.store.add(self._k(key), amount)
    def check(self, keys): return self.store.check([self._k(k) for k in ([keys] if isinstance(keys, str) else keys)])
    def wait(self, keys, timeout=None): self.store.wait([self._k(k) for k in ([keys] if isinstance(keys, str) else keys)], timeout)
    def delete_key(self, key): return self.store.delete_key(self._k(key))
    def set_obj(self, key, obj): self.store.set_obj(self._k(key), obj)
    def get_obj(self, key, timeout=None): return self.store.get_obj(self._k(key), timeout)
    def barrier(self, name, rank, world_size, timeout=None): self.store.barrier(self._k(name), rank, world_size, timeout)

    @property
    def client_handle(self): return self.store.client_handle
    @property
    def host(self): return self.store.host


def rendezvous(init_method: Optional[str], rank: int, world_size: int, timeout: float = DEFAULT_TIMEOUT_S):
    """Resolve (store, rank, world_size) from ``env://`` or ``tcp://host:port``.

    ``env://`` reads MASTER_ADDR / MASTER_PORT (and RANK / WORLD_SIZE when not passed), exactly the
    variables the reference sets before ``init_process_group`` (ref:main.py:48-49).
    """
    init_method = init_method or "env://"
    if init_method.startswith("env://"):
        host = os.environ.get("MASTER_ADDR", "127.0.0.1")
        port = int(os.environ.get("MASTER_PORT", DEFAULT_PORT))
        if rank < 0:
            rank = int(os.environ["RANK"])
        if world_size < 0:
            world_size = int(os.environ["WORLD_SIZE"])
    elif init_method.startswith("tcp://"):
        hostport = init_method[len("tcp://"):].split("?")[0]
        host, port_s = hostport.rsplit(":", 1)
        port = int(port_s)
        if rank < 0 or world_size < 0:
            raise ValueError("tcp:// rendezvous needs explicit rank and world_size")
    else:
        raise ValueError(f"unsupported init_method {init_method!r} (use env:// or tcp://host:port)")
    if host == "localhost":
        host = "127.0.0.1"
    if init_method.startswith("env://") and os.environ.get("TORCHELASTIC_USE_AGENT_STORE", "").lower() == "true":
        return _rendezvous_behind_agent(host, port, rank, world_size, timeout), rank, world_size
    store = TCPStore(host, port, world_size, is_master=(rank == 0), timeout=timeout)
    return store, rank, world_size


_agent_inits = 0


def _free_port(host: str) -> int:
    import socket
    with socket.socket() as s:
        s.bind((host, 0))
        return s.getsockname()[1]


def _rendezvous_behind_agent(host: str, port: int, rank: int, world_size: int, timeout: float) -> "TCPStore":
    """``torchrun --master-addr A --master-port P`` (static rendezvous) keeps ITS OWN store listening on P and tells the
    workers so with TORCHELASTIC_USE_AGENT_STORE=True: MASTER_PORT is taken.  Rank 0 then opens this package's store on
    a free port and publishes the number through the launcher's store (the one piece of torch.distributed plumbing
    used here); the other ranks read it there."""
    global _agent_inits
    from datetime import timedelta

    from torch.distributed import TCPStore as _LauncherStore
    agent = _LauncherStore(host, port, world_size, False, timedelta(seconds=timeout), wait_for_workers=False)
    key = "dcp_b200/store_port/%s/%d" % (os.environ.get("TORCHELASTIC_RESTART_COUNT", "0"), _agent_inits)
    _agent_inits += 1
    if rank == 0:
        last = None
        for _ in range(16):                       # another process may grab the port between the probe and the listen
            own = _free_port(host)
            try:
                store = TCPStore(host, own, world_size, is_master=True, timeout=timeout)
            except RuntimeError as e:
                last = e
                continue
            agent.set(key, str(own))
            return store
        raise RuntimeError(f"rendezvous behind the launcher's store failed: {last}")
    own = int(agent.get(key).decode())
    return TCPStore(host, own, world_size, is_master=False, timeout=timeout)
