"""Fused flat-arena optimizers (K6).

``FusedAdam`` / ``FusedSGD`` update the whole model with ONE kernel launch over the flat
parameter arena instead of the per-tensor Python loop of ``torch.optim.Adam`` /
``torch.optim.SGD`` that the reference uses (src/motion/trainer/base.py:42-43,117;
src/example/example_ddp.py:54,82).  The math is torch's (Adam: no amsgrad, bias-corrected,
``eps`` added after the sqrt; SGD: optional momentum, L2 weight decay), and
``state_dict()`` / ``load_state_dict()`` speak the stock ``torch.optim`` format, so a
checkpoint written here loads into ``torch.optim.Adam`` unchanged and vice versa
(SURVEY §5.4).

On CPU tensors the same update runs as a handful of flat torch ops (plumbing / oracle path).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import torch

from .. import _native
from .._native import OptimDesc
from ..utils.arena import ParamArena


class _FlatOptimizer(torch.optim.Optimizer):
    kind = 0

    def __init__(self, arena: ParamArena, defaults: dict):
        self.arena = arena
        super().__init__(list(arena.module.parameters()), defaults)
        P = arena.numel
        self.exp_avg = torch.zeros(P, dtype=torch.float32, device=arena.device)
        self.exp_avg_sq = torch.zeros(P, dtype=torch.float32, device=arena.device)
        self.step_count = 0
        self.grad_scale = 1.0

    # -- the launch --
    def desc(self, step: Optional[int] = None, grad_scale: Optional[float] = None) -> OptimDesc:
        g = self.param_groups[0]
        o = OptimDesc()
        o.kind = self.kind
        o.step = self.step_count if step is None else step
        o.lr = g["lr"]
        b = g.get("betas", (0.0, 0.0))
        o.beta1, o.beta2 = b
        o.eps = g.get("eps", 0.0)
        o.weight_decay = g.get("weight_decay", 0.0)
        o.momentum = g.get("momentum", 0.0)
        o.grad_scale = self.grad_scale if grad_scale is None else grad_scale
        return o

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self.step_count += 1
        a = self.arena
        if a.flat.is_cuda:
            lib = _native.lib()
            o = self.desc()
            _native.check(lib.b200rnn_fused_update(
                a.grad.data_ptr(), 1, a.numel, a.flat.data_ptr(), self.exp_avg.data_ptr(),
                self.exp_avg_sq.data_ptr(), C.byref(o), None, None, _native.stream_ptr(a.device)), "fused_update")
            _native.count_launch()
        else:
            self._step_torch(a.grad[:a.numel] * self.grad_scale if self.grad_scale != 1.0 else a.grad[:a.numel])
        return loss

    def _step_torch(self, g: torch.Tensor) -> None:
        raise NotImplementedError

    def zero_grad(self, set_to_none: bool = False):  # grads are arena views: never drop them
        self.arena.zero_grad()

    # -- stock torch.optim state_dict format --
    def _per_param_state(self, i: int, off: int, k: int, shp) -> dict:
        raise NotImplementedError

    def state_dict(self):
        groups = []
        for g in self.param_groups:
            gg = {k: v for k, v in g.items() if k != "params"}
            gg["params"] = list(range(len(g["params"])))
            groups.append(gg)
        state = {}
        if self.step_count > 0:
            for i, (name, (off, k, shp)) in enumerate(self.arena.index.items()):
                st = self._per_param_state(i, off, k, shp)
                if st:
                    state[i] = st
        return {"state": state, "param_groups": groups}

    def load_state_dict(self, sd):
        groups = sd["param_groups"]
        if len(groups) != 1:
            raise ValueError("flat optimizers hold exactly one param group")
        g = self.param_groups[0]
        for k, v in groups[0].items():
            if k != "params":
                g[k] = v
        self._load_state(sd.get("state", {}))

    def _load_state(self, state: dict) -> None:
        raise NotImplementedError


class FusedAdam(_FlatOptimizer):
    kind = 0

    def __init__(self, arena: ParamArena, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0):
        defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay, amsgrad=False,
                        maximize=False, foreach=None, capturable=False, differentiable=False, fused=None,
                        decoupled_weight_decay=False)
        super().__init__(arena, defaults)

    def _step_torch(self, g):
        grp = self.param_groups[0]
        b1, b2 = grp["betas"]
        p = self.arena.flat[:self.arena.numel]
        if grp["weight_decay"]:
            g = g + grp["weight_decay"] * p
        self.exp_avg.mul_(b1).add_(g, alpha=1 - b1)
        self.exp_avg_sq.mul_(b2).addcmul_(g, g, value=1 - b2)
        bc1 = 1 - b1 ** self.step_count
        bc2 = 1 - b2 ** self.step_count
        denom = (self.exp_avg_sq.sqrt() / math.sqrt(bc2)).add_(grp["eps"])
        p.addcdiv_(self.exp_avg, denom, value=-grp["lr"] / bc1)

    def _per_param_state(self, i, off, k, shp):
        return {"step": torch.tensor(float(self.step_count)),
                "exp_avg": self.exp_avg[off:off + k].view(shp).clone(),
                "exp_avg_sq": self.exp_avg_sq[off:off + k].view(shp).clone()}

    def _load_state(self, state):
        steps = set()
        for i, (name, (off, k, shp)) in enumerate(self.arena.index.items()):
            st = state.get(i, state.get(str(i)))
            if st is None:
                continue
            self.exp_avg[off:off + k].copy_(st["exp_avg"].reshape(-1))
            self.exp_avg_sq[off:off + k].copy_(st["exp_avg_sq"].reshape(-1))
            steps.add(int(float(st["step"])))
        if len(steps) > 1:
            raise ValueError(f"per-parameter step counts differ ({sorted(steps)}); a flat optimizer needs one")
        self.step_count = steps.pop() if steps else 0


class FusedSGD(_FlatOptimizer):
    kind = 1

    def __init__(self, arena: ParamArena, lr: float = 1e-3, momentum: float = 0.0, weight_decay: float = 0.0):
        defaults = dict(lr=lr, momentum=momentum, dampening=0, weight_decay=weight_decay, nesterov=False,
                        maximize=False, foreach=None, differentiable=False, fused=None)
        super().__init__(arena, defaults)

    def _step_torch(self, g):
        grp = self.param_groups[0]
        p = self.arena.flat[:self.arena.numel]
        if grp["weight_decay"]:
            g = g + grp["weight_decay"] * p
        if grp["momentum"]:
            self.exp_avg.mul_(grp["momentum"]).add_(g)
            g = self.exp_avg
        p.add_(g, alpha=-grp["lr"])

    def _per_param_state(self, i, off, k, shp):
        if not self.param_groups[0]["momentum"]:
            return {}
        return {"momentum_buffer": self.exp_avg[off:off + k].view(shp).clone()}

    def _load_state(self, state):
        for i, (name, (off, k, shp)) in enumerate(self.arena.index.items()):
            st = state.get(i, state.get(str(i)))
            if st and st.get("momentum_buffer") is not None:
                self.exp_avg[off:off + k].copy_(st["momentum_buffer"].reshape(-1))
