"""R1: the C++ bootstrap store (csrc/store.cpp) through its C ABI. CPU only."""
import threading
import time

import pytest

from pdcc_b200 import _native as N


@pytest.fixture()
def server():
    s = N.Store("127.0.0.1", 0, serve=True)
    yield s
    s.close()


def test_set_get_add_check_delete(server):
    c = N.Store("127.0.0.1", server.port)
    server.set("a", b"hello")
    assert c.get("a") == b"hello"
    assert c.add("n", 5) == 5 and server.add("n", -2) == 3
    assert c.check("a") and not c.check("zzz")
    c.delete("a")
    assert not server.check("a")
    c.set("empty", b"")
    assert server.get("empty") == b""
    big = bytes(range(256)) * 1024
    c.set("big", big)
    assert server.get("big") == big
    c.close()


def test_get_blocks_until_set(server):
    c = N.Store("127.0.0.1", server.port)
    t0 = time.monotonic()
    threading.Timer(0.3, lambda: server.set("late", b"x")).start()
    assert c.get("late", timeout_ms=5000) == b"x"
    assert time.monotonic() - t0 >= 0.25
    c.close()


def test_get_times_out(server):
    with pytest.raises(N.PdccError) as e:
        server.get("never", timeout_ms=100)
    assert e.value.code == N.ETIMEOUT and "never" in str(e.value)


def test_barrier_many_clients(server):
    n = 8
    order = []

    def worker(i):
        c = N.Store("127.0.0.1", server.port)
        time.sleep(0.02 * i)
        c.barrier("b0", n, 10000)
        order.append(time.monotonic())
        c.close()

    th = [threading.Thread(target=worker, args=(i,)) for i in range(n)]
    t0 = time.monotonic()
    [t.start() for t in th]
    [t.join() for t in th]
    assert len(order) == n and min(order) - t0 >= 0.02 * (n - 1) - 0.01


def test_connect_timeout():
    with pytest.raises(N.PdccError) as e:
        N.Store("127.0.0.1", 1, timeout_ms=200)
    assert e.value.code == N.ETIMEOUT
