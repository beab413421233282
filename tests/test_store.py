import threading
import time

import pytest

from distributed_compute_pytorch_b200.store import PrefixStore, StoreTimeout, TCPStore


def test_set_get_add_check_delete(free_port):
    master = TCPStore("127.0.0.1", free_port, 2, is_master=True)
    client = TCPStore("127.0.0.1", free_port, 2, is_master=False)
    master.set("a", b"hello")
    assert client.get("a") == b"hello"
    assert client.add("ctr", 3) == 3
    assert master.add("ctr", 4) == 7
    assert client.check(["a", "ctr"]) and not client.check("nope")
    assert client.delete_key("a") and not client.check("a")
    big = bytes(range(256)) * 4096  # 1 MiB value
    client.set("big", big)
    assert master.get("big") == big
    client.set_obj("obj", {"x": [1, 2, 3]})
    assert master.get_obj("obj") == {"x": [1, 2, 3]}
    client.close(); master.close()


def test_get_blocks_until_set(free_port):
    master = TCPStore("127.0.0.1", free_port, 2, is_master=True)
    client = TCPStore("127.0.0.1", free_port, 2)
    got = {}
    t = threading.Thread(target=lambda: got.setdefault("v", client.get("late", timeout=10)))
    t.start()
    time.sleep(0.2)
    assert "v" not in got
    master.set("late", b"42")
    t.join(5)
    assert got["v"] == b"42"
    client.close(); master.close()


def test_get_timeout_and_recovery(free_port):
    master = TCPStore("127.0.0.1", free_port, 1, is_master=True)
    with pytest.raises(StoreTimeout):
        master.get("never", timeout=0.2)
    master.set("k", b"v")  # the client reconnected after the timeout
    assert master.get("k") == b"v"
    master.close()


def test_prefix_store_and_barrier(free_port):
    master = TCPStore("127.0.0.1", free_port, 2, is_master=True)
    other = TCPStore("127.0.0.1", free_port, 2)
    a, b = PrefixStore("pg0", master), PrefixStore("pg1", master)
    a.set("k", b"1"); b.set("k", b"2")
    assert a.get("k") == b"1" and b.get("k") == b"2"
    done = []
    t = threading.Thread(target=lambda: (PrefixStore("pg0", other).barrier("bar", 1, 2, 10), done.append(1)))
    t.start()
    time.sleep(0.1)
    assert not done
    a.barrier("bar", 0, 2, 10)
    t.join(5)
    assert done == [1]
    other.close(); master.close()


def test_port_in_use(free_port):
    m1 = TCPStore("127.0.0.1", free_port, 1, is_master=True)
    with pytest.raises(RuntimeError):
        TCPStore("127.0.0.1", free_port, 1, is_master=True)
    m1.close()
