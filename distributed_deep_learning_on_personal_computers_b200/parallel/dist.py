"""Rank / world bootstrap and the rank-0 state broadcast.

Reference behaviour being covered (``R`` = the reference script):

* identity: the reference looks its rank up from the host name in a hard-coded table and makes
  rank 0 the server (``R:226-233``, ``R:248-249``, ``R:690``, ``R:792``).  Here rank and world size
  come from the config / the ``torchrun`` environment; rank 0 is still the coordinator (it owns
  the run log, the checkpoints and the sample images).
* rendezvous: ``create_server`` accepts exactly ``N_conn`` connections, ``create_worker``
  connects once with no retry and no timeout (``R:173-206``).  Here ``init`` blocks until all
  ``world_size`` ranks joined, with a timeout, through ``torch.distributed`` (NCCL on GPUs,
  gloo on CPU).
* initial state: the server pickles the live ``[model, optimizer, loss]`` objects and sends them
  to every worker (``R:560-565``).  Here rank 0 broadcasts the flat parameter buffer and the BN
  buffers; the optimizer and the loss are rebuilt locally from the config (they are stateless
  at step 0), which is equivalent and does not ship code over the wire.
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass
class DistContext:
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    device: torch.device = torch.device("cpu")
    backend: str = "none"
    initialized_here: bool = False
    star: object = None              # backend "star": the TCP endpoint (StarServer on rank 0, StarWorker elsewhere)

    @property
    def is_coordinator(self) -> bool:
        return self.rank == 0

    def barrier(self) -> None:
        if self.world_size > 1:
            if self.backend == "star":
                # the reference has no barrier; its blocking recv is one (R:215-220).  Same thing, explicit.
                if self.rank == 0:
                    self.star.gather()
                    self.star.broadcast({"barrier": 1})
                else:
                    self.star.send({"barrier": 1})
                    self.star.recv()
            elif self.backend == "nccl":
                dist.barrier(device_ids=[self.device.index])
            else:
                dist.barrier()

    def shutdown(self) -> None:
        if self.star is not None:
            self.star.close()
            self.star = None
        if self.initialized_here and dist.is_initialized():
            dist.destroy_process_group()
            self.initialized_here = False


def pick_device(cfg) -> torch.device:
    want = cfg.device
    if want == "auto":
        want = "cuda" if torch.cuda.is_available() else "cpu"
    if want == "cuda":
        idx = cfg.local_rank % max(torch.cuda.device_count(), 1)
        torch.cuda.set_device(idx)
        return torch.device("cuda", idx)
    return torch.device("cpu")


def init(cfg) -> DistContext:
    """Join the job described by ``cfg`` (blocks until every rank is present)."""
    device = pick_device(cfg)
    if cfg.world_size <= 1:
        return DistContext(0, 1, cfg.local_rank, device, "none", False)
    if cfg.comm == "star":
        # The reference's own bootstrap: bare TCP sockets, rank 0 listens and accepts world-1 connections
        # (R:173-206).  No torch.distributed process group exists on this path: rendezvous, the initial
        # state broadcast, the gradient rounds and the barriers all go over the star.
        return DistContext(cfg.rank, cfg.world_size, cfg.local_rank, device, "star", False, open_star(cfg))
    backend = cfg.backend
    if backend == "auto":
        backend = "nccl" if device.type == "cuda" else "gloo"
    started = False
    if not dist.is_initialized():
        kwargs = dict(backend=backend, rank=cfg.rank, world_size=cfg.world_size,
                      timeout=datetime.timedelta(seconds=cfg.timeout_s))
        if all(k in os.environ for k in ("MASTER_ADDR", "MASTER_PORT", "RANK", "WORLD_SIZE")):
            kwargs["init_method"] = "env://"
        else:
            kwargs["init_method"] = f"tcp://{cfg.master_addr}:{cfg.master_port}"
        if backend == "nccl":
            kwargs["device_id"] = device
        dist.init_process_group(**kwargs)
        started = True
    return DistContext(dist.get_rank(), dist.get_world_size(), cfg.local_rank, device,
                       dist.get_backend(), started)


def open_star(cfg):
    """Rank 0 listens on ``master_addr:master_port + 1`` and waits for every worker; workers connect with
    retries (R:173-206, with the timeouts the reference lacks)."""
    from . import star
    kw = dict(compress=cfg.compress, level=cfg.compress_level, threads=cfg.compress_threads,
              block=cfg.compress_block, strategy=cfg.compress_strategy)
    port = cfg.master_port + 1
    if cfg.rank == 0:
        srv = star.StarServer(cfg.master_addr, port, cfg.world_size - 1, cfg.timeout_s, **kw)
        srv.accept_all()
        return srv
    return star.StarWorker(cfg.master_addr, port, cfg.rank, cfg.timeout_s, **kw)


@torch.no_grad()
def broadcast_model_state(model, ctx: DistContext, src: int = 0) -> None:
    """Make every replica start from rank ``src``'s parameters and BN buffers (R:560-565)."""
    if ctx.world_size <= 1:
        return
    if ctx.backend == "star":
        # tensors, not pickled modules: rank 0 packs the flat parameter buffer and the BN buffers once
        flat = model.flat
        if ctx.rank == 0:
            ctx.star.broadcast({"param": flat["param"].detach().cpu(),
                                "buffers": [b.detach().cpu() for b in model.buffers()]})
        else:
            msg = ctx.star.recv()
            flat["param"].copy_(msg["param"].to(flat["param"].device).view_as(flat["param"]))
            for b, v in zip(model.buffers(), msg["buffers"]):
                b.copy_(v.to(b.device).view_as(b))
        return
    flat = getattr(model, "_flat", None)
    if flat:
        dist.broadcast(flat["param"], src=src)
    else:
        for p in model.parameters():
            dist.broadcast(p.data, src=src)
    for b in model.buffers():
        dist.broadcast(b.data, src=src)


def all_reduce_max_time(ms: float, ctx: DistContext) -> float:
    """Max over ranks of a device-measured duration (multi-GPU numbers are max-over-ranks)."""
    if ctx.world_size <= 1 or ctx.backend == "star":
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=ctx.device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
