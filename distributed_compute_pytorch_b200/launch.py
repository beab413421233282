"""Multi-process launcher: ``spawn(fn, args, nprocs, join)`` / ``launch(...)``.

Same calling convention as the ``mp.spawn(proc, args=(world_size, opt, use_cuda), nprocs=world_size,
join=True)`` line of the reference (ref:main.py:150): one child per rank, ``fn(rank, *args)``, and the
parent waits.  Failure semantics match what users of that call rely on: the first child that raises
(or dies) causes every sibling to be terminated and the parent to raise with the child's traceback.
"""
from __future__ import annotations

import multiprocessing as mp
import os
import signal
import sys
import tempfile
import time
import traceback
from typing import Callable, Optional, Sequence


class ProcessRaisedException(RuntimeError):
    def __init__(self, msg: str, error_index: int, error_pid: int):
        super().__init__(msg)
        self.error_index, self.error_pid = error_index, error_pid


class ProcessExitedException(RuntimeError):
    def __init__(self, msg: str, error_index: int, error_pid: int, exit_code: int, signal_name: Optional[str] = None):
        super().__init__(msg)
        self.error_index, self.error_pid, self.exit_code, self.signal_name = error_index, error_pid, exit_code, signal_name


def _child(fn: Callable, rank: int, args: Sequence, err_path: str, env: dict) -> None:
    os.environ.update(env)
    try:
        fn(rank, *args)
    except KeyboardInterrupt:
        pass
    except Exception:
        with open(err_path, "w") as f:
            f.write(traceback.format_exc())
        sys.exit(1)


class ProcessContext:
    def __init__(self, processes, err_files):
        self.processes, self.err_files = processes, err_files

    def pids(self):
        return [p.pid for p in self.processes]

    def _terminate_all(self) -> None:
        for p in self.processes:
            if p.is_alive():
                p.terminate()
        deadline = time.time() + 10
        for p in self.processes:
            p.join(max(0.0, deadline - time.time()))
        for p in self.processes:
            if p.is_alive():
                p.kill()
                p.join()

    def join(self, timeout: Optional[float] = None) -> bool:
        """Returns True once every child exited cleanly; raises as soon as one fails."""
        deadline = None if timeout is None else time.time() + timeout
        while True:
            alive = False
            for i, p in enumerate(self.processes):
                if p.is_alive():
                    alive = True
                    continue
                if p.exitcode not in (0, None):
                    self._terminate_all()
                    msg = ""
                    if os.path.exists(self.err_files[i]):
                        with open(self.err_files[i]) as f:
                            msg = f.read()
                    self._cleanup()
                    if msg:
                        raise ProcessRaisedException(f"\n\n-- Process {i} terminated with the following error:\n{msg}", i, p.pid)
                    if p.exitcode < 0:
                        try:
                            name = signal.Signals(-p.exitcode).name
                        except ValueError:
                            name = f"<signal {-p.exitcode}>"
                        raise ProcessExitedException(f"process {i} terminated with signal {name}", i, p.pid, p.exitcode, name)
                    raise ProcessExitedException(f"process {i} terminated with exit code {p.exitcode}", i, p.pid, p.exitcode)
            if not alive:
                self._cleanup()
                return True
            if deadline is not None and time.time() > deadline:
                return False
            time.sleep(0.02)

    def _cleanup(self) -> None:
        for f in self.err_files:
            try:
                os.unlink(f)
            except OSError:
                pass


def start_processes(fn: Callable, args: Sequence = (), nprocs: int = 1, join: bool = True, daemon: bool = False,
                    start_method: str = "spawn", env: Optional[dict] = None):
    ctx = mp.get_context(start_method)
    procs, errs = [], []
    for rank in range(nprocs):
        fd, path = tempfile.mkstemp(prefix=f"dcp_err_{rank}_", suffix=".txt")
        os.close(fd)
        os.unlink(path)
        child_env = dict(env or {})
        p = ctx.Process(target=_child, args=(fn, rank, tuple(args), path, child_env), daemon=daemon)
        p.start()
        procs.append(p)
        errs.append(path)
    context = ProcessContext(procs, errs)
    if not join:
        return context
    while not context.join():
        pass
    return None


def spawn(fn: Callable, args: Sequence = (), nprocs: int = 1, join: bool = True, daemon: bool = False,
          start_method: str = "spawn"):
    """``mp.spawn``-compatible entry point (ref:main.py:150)."""
    return start_processes(fn, args, nprocs, join, daemon, start_method)


def launch(fn: Callable, nprocs: int, args: Sequence = (), master_addr: str = "127.0.0.1", master_port: Optional[int] = None,
           join: bool = True):
    """Spawn ``nprocs`` ranks with MASTER_ADDR/MASTER_PORT/WORLD_SIZE exported (RANK/LOCAL_RANK are set
    per child by ``fn``'s first argument: ``fn(rank, *args)``)."""
    if master_port is None:
        import socket
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            master_port = s.getsockname()[1]
    env = {"MASTER_ADDR": master_addr, "MASTER_PORT": str(master_port), "WORLD_SIZE": str(nprocs)}
    return start_processes(fn, args, nprocs, join, env=env)
