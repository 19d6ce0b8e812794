"""The file-descriptor exchange of the peer plane (csrc/mem.cu: fedb_fd_serve / fedb_fd_fetch) through the C ABI.  It carries
the POSIX handles of the GPU allocations between worker processes (SCM_RIGHTS over a unix socket); the mechanism itself needs
no GPU, so it is checked here with ordinary descriptors: the bytes behind a fetched descriptor are the server's, the socket is
owner-only, a client that never acknowledges cannot wedge the server, and a missing server is a bounded error."""
import ctypes as C
import os
import socket
import stat
import threading
import time

import pytest

from fedb200.ops import _lib


@pytest.fixture(scope="module")
def L():
    if not os.path.exists(_lib.LIB_PATH):
        pytest.skip("libfedb200.so is not built")
    return _lib.lib()


def _serve(L, path, fds, n_clients, timeout_ms, out):
    arr = (C.c_int * len(fds))(*fds)
    out.append(L.fedb_fd_serve(path.encode(), arr, len(fds), n_clients, timeout_ms))


def test_descriptors_travel_and_the_socket_is_owner_only(L, tmp_path):
    path = str(tmp_path / "x.sock")
    r, w = os.pipe()
    os.write(w, b"arena-bytes")
    out = []
    th = threading.Thread(target=_serve, args=(L, path, [r, w], 2, 10_000, out))
    th.start()
    got = (C.c_int * 2)()
    assert L.fedb_fd_fetch(path.encode(), got, 2, 5000) == 0            # waits for the socket to appear
    mode = stat.S_IMODE(os.stat(path).st_mode)
    assert mode & 0o077 == 0, oct(mode)                                  # nobody but the owner may connect
    assert os.read(got[0], 64) == b"arena-bytes"                         # the SAME pipe, seen through the received descriptor
    os.write(got[1], b"second")
    assert os.read(r, 64) == b"second"
    got2 = (C.c_int * 2)()
    assert L.fedb_fd_fetch(path.encode(), got2, 2, 5000) == 0
    th.join(10)
    assert out == [0] and not os.path.exists(path)                       # served its two clients, cleaned up
    for fd in (r, w, got[0], got[1], got2[0], got2[1]):
        os.close(fd)


def test_a_client_that_never_acknowledges_does_not_wedge_the_server(L, tmp_path):
    path = str(tmp_path / "y.sock")
    r, w = os.pipe()
    out = []
    th = threading.Thread(target=_serve, args=(L, path, [r], 2, 20_000, out))
    th.start()
    deadline = time.time() + 5
    while not os.path.exists(path) and time.time() < deadline:
        time.sleep(0.01)
    mute = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    mute.connect(path)                                                   # takes the descriptors' message, never answers
    t0 = time.time()
    got = (C.c_int * 1)()
    assert L.fedb_fd_fetch(path.encode(), got, 1, 15_000) == 0           # the well-behaved client is still served
    assert time.time() - t0 < 12                                         # after the bounded wait for the mute one's ack
    th.join(10)
    assert out == [0]
    mute.close()
    for fd in (r, w, got[0]):
        os.close(fd)


def test_fetch_from_nobody_times_out_with_a_message(L, tmp_path):
    got = (C.c_int * 1)()
    t0 = time.time()
    rc = L.fedb_fd_fetch(str(tmp_path / "nobody.sock").encode(), got, 1, 300)
    assert rc != 0 and 0.25 < time.time() - t0 < 3
    assert b"timeout" in L.fedb_last_error()
