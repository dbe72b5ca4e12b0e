"""Fused optimizers over the flat parameter / gradient storage.

The reference builds ``torch.optim.Adam(model.parameters())`` with torch defaults (``R:704``) and
calls ``optimizer.step(); optimizer.zero_grad()`` after every gradient exchange (``R:437-438``,
``R:552-553``): 100 parameter tensors, i.e. (without ``foreach``) hundreds of small launches.
Here one kernel (``csrc/elem/optim.cu``) walks the single flat buffer: it applies the pending
gradient scale (1/world, 1/K), updates ``m``/``v``/``p`` in fp32, refreshes the bf16 weight copy
that the tensor-core kernels read, and zeroes the gradient for the next accumulation window --
one launch instead of ``step`` + ``zero_grad`` + a cast pass.

``state_dict()`` is laid out like ``torch.optim.Adam.state_dict()`` (per-parameter ``step``,
``exp_avg``, ``exp_avg_sq`` in logical shape), so the optimizer state can be loaded into a stock
``torch.optim.Adam`` built on the reference's model and back.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from .ops import functional as Fn


class _FlatOptimizer:
    def __init__(self, model, lr: float, weight_decay: float):
        flat = model.flat
        self.model = model
        self.p: torch.Tensor = flat["param"]
        self.g: torch.Tensor = flat["grad"]
        self.offsets = flat["offsets"]
        self.lr = lr
        self.weight_decay = weight_decay
        self.step_count = 0
        self.p_lowp: Optional[torch.Tensor] = None

    def attach_lowp(self, p_lowp: Optional[torch.Tensor]) -> None:
        """Have ``step`` also refresh this low-precision copy of the parameters."""
        self.p_lowp = p_lowp if (p_lowp is not None and p_lowp is not self.p) else None

    def rebind_grad(self, g: torch.Tensor) -> None:
        self.g = g

    def zero_grad(self) -> None:
        self.g.zero_()

    def _per_param(self, flat_state: torch.Tensor, name: str, p: torch.Tensor) -> torch.Tensor:
        off, n = self.offsets[name]
        from .models.unet import UNet
        is_t = p.dim() == 4 and name.endswith("up_sample.weight")
        view = UNet._internal_view(flat_state, off, p, is_t)
        return view.detach().cpu().contiguous().clone()

    def _load_per_param(self, flat_state: torch.Tensor, name: str, p: torch.Tensor,
                        value: torch.Tensor) -> None:
        from .models.unet import UNet
        off, _ = self.offsets[name]
        is_t = p.dim() == 4 and name.endswith("up_sample.weight")
        UNet._internal_view(flat_state, off, p, is_t).copy_(value.to(flat_state.device))


class FlatAdam(_FlatOptimizer):
    def __init__(self, model, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        super().__init__(model, lr, weight_decay)
        self.betas = tuple(betas)
        self.eps = eps
        self.m = torch.zeros_like(self.p)
        self.v = torch.zeros_like(self.p)

    @torch.no_grad()
    def step(self, grad_scale: float = 1.0, zero_grad: bool = True) -> None:
        self.step_count += 1
        Fn.fused_adam(self.p, self.g, self.m, self.v, self.step_count, self.lr, self.betas[0],
                      self.betas[1], self.eps, self.weight_decay, grad_scale, self.p_lowp,
                      zero_grad)

    def state_dict(self) -> Dict:
        state = {}
        names = []
        for i, (name, p) in enumerate(self.model.named_parameters()):
            names.append(name)
            state[i] = {"step": torch.tensor(float(self.step_count)),
                        "exp_avg": self._per_param(self.m, name, p),
                        "exp_avg_sq": self._per_param(self.v, name, p)}
        group = {"lr": self.lr, "betas": self.betas, "eps": self.eps,
                 "weight_decay": self.weight_decay, "amsgrad": False, "maximize": False,
                 "foreach": None, "capturable": False, "differentiable": False, "fused": None,
                 "decoupled_weight_decay": False, "params": list(range(len(names)))}
        return {"state": state if self.step_count else {}, "param_groups": [group],
                "kind": "adam", "param_names": names}

    def load_state_dict(self, sd: Dict) -> None:
        g = sd["param_groups"][0]
        self.lr, self.betas, self.eps = g["lr"], tuple(g["betas"]), g["eps"]
        self.weight_decay = g.get("weight_decay", 0.0)
        self.m.zero_()
        self.v.zero_()
        self.step_count = 0
        for i, (name, p) in enumerate(self.model.named_parameters()):
            st = sd["state"].get(i)
            if st is None:
                continue
            self.step_count = int(float(st["step"]))
            self._load_per_param(self.m, name, p, st["exp_avg"])
            self._load_per_param(self.v, name, p, st["exp_avg_sq"])


class FlatSGD(_FlatOptimizer):
    def __init__(self, model, lr=1e-2, momentum=0.0, weight_decay=0.0):
        super().__init__(model, lr, weight_decay)
        self.momentum = momentum
        self.buf = torch.zeros_like(self.p) if momentum else None

    @torch.no_grad()
    def step(self, grad_scale: float = 1.0, zero_grad: bool = True) -> None:
        self.step_count += 1
        Fn.fused_sgd(self.p, self.g, self.buf, self.lr, self.momentum, self.weight_decay,
                     grad_scale, self.p_lowp, zero_grad, first_step=(self.step_count == 1))

    def state_dict(self) -> Dict:
        state = {}
        names = [n for n, _ in self.model.named_parameters()]
        if self.buf is not None and self.step_count:
            for i, (name, p) in enumerate(self.model.named_parameters()):
                state[i] = {"momentum_buffer": self._per_param(self.buf, name, p)}
        group = {"lr": self.lr, "momentum": self.momentum, "dampening": 0,
                 "weight_decay": self.weight_decay, "nesterov": False, "maximize": False,
                 "foreach": None, "differentiable": False, "fused": None,
                 "params": list(range(len(names)))}
        return {"state": state, "param_groups": [group], "kind": "sgd", "param_names": names,
                "step_count": self.step_count}

    def load_state_dict(self, sd: Dict) -> None:
        g = sd["param_groups"][0]
        self.lr, self.momentum = g["lr"], g["momentum"]
        self.weight_decay = g.get("weight_decay", 0.0)
        self.step_count = int(sd.get("step_count", 0))
        if self.buf is not None:
            self.buf.zero_()
            for i, (name, p) in enumerate(self.model.named_parameters()):
                st = sd["state"].get(i)
                if st is not None:
                    self._load_per_param(self.buf, name, p, st["momentum_buffer"])


def build_optimizer(model, cfg):
    if cfg.optimizer == "adam":
        return FlatAdam(model, cfg.lr, cfg.betas, cfg.eps, cfg.weight_decay)
    if cfg.optimizer == "sgd":
        return FlatSGD(model, cfg.lr, cfg.momentum, cfg.weight_decay)
    raise ValueError(f"unknown optimizer {cfg.optimizer!r}")
