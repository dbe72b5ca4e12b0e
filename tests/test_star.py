"""TCP-star transport: framing, deflate stage, rendezvous, and the fault paths that crash the
reference (peer EOF, timeout, short read -- SURVEY 5.3 / Q14)."""
import gzip
import socket
import struct
import threading

import pytest
import torch

from distributed_deep_learning_on_personal_computers_b200.parallel import star


def test_deflate_is_multimember_gzip_and_roundtrips():
    data = bytes(range(256)) * 20_000
    z = star.deflate(data, level=1, threads=4, block=1_000_000)
    assert gzip.decompress(z) == data            # any gzip reader (e.g. mgzip's peer) can read it
    assert star.inflate(z) == data
    assert star.inflate(star.deflate(b"tiny")) == b"tiny"


def test_pack_roundtrip_tensors():
    msg = {"max": 0.25, "q": torch.arange(10, dtype=torch.float16),
           "b": torch.randn(3, 4).to(torch.bfloat16), "i": torch.tensor([1, -2], dtype=torch.int8)}
    for comp in (True, False):
        out = star.unpack(star.pack(msg, compress=comp), compressed=comp)
        assert out["max"] == 0.25
        for k in ("q", "b", "i"):
            assert out[k].dtype == msg[k].dtype and torch.equal(out[k], msg[k])


def _pair():
    a, b = socket.socketpair()
    a.settimeout(2)
    b.settimeout(2)
    return a, b


def test_frame_is_big_endian_length_prefixed():
    a, b = _pair()
    star.send_frame(a, b"hello")
    raw = b.recv(64)
    assert raw[:4] == struct.pack(">I", 5) and raw[4:] == b"hello"


def test_peer_eof_raises_peerlost():
    a, b = _pair()
    a.sendall(struct.pack(">I", 100) + b"short")
    a.close()
    with pytest.raises(star.PeerLost):
        star.recv_frame(b)


def test_timeout_raises_peerlost():
    a, b = _pair()
    b.settimeout(0.2)
    with pytest.raises(star.PeerLost):
        star.recv_frame(b)


def test_server_worker_gather_broadcast():
    srv = star.StarServer("127.0.0.1", 0, n_workers=2, timeout=10)
    got = {}

    def worker(rank):
        w = star.StarWorker("127.0.0.1", srv.port, rank, timeout=10)
        w.send({"rank": rank, "t": torch.full((4,), float(rank))})
        got[rank] = w.recv()
        w.close()

    ths = [threading.Thread(target=worker, args=(r,)) for r in (2, 1)]
    for t in ths:
        t.start()
    srv.accept_all()
    msgs = srv.gather()
    assert [m["rank"] for m in msgs] == [1, 2]          # rank order, not accept order
    srv.broadcast({"sum": sum(m["t"] for m in msgs)})
    for t in ths:
        t.join()
    assert torch.equal(got[1]["sum"], torch.full((4,), 3.0))
    srv.close()


def test_worker_gives_up_when_no_server():
    with pytest.raises(star.PeerLost):
        star.StarWorker("127.0.0.1", 1, 1, timeout=0.2, connect_retries=3, retry_delay=0.01)


def test_missing_worker_times_out():
    srv = star.StarServer("127.0.0.1", 0, n_workers=1, timeout=0.3)
    with pytest.raises(star.PeerLost):
        srv.accept_all()
    srv.close()
