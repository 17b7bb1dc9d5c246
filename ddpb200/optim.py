"""SGD with momentum and weight decay, fused over the flat parameter arena.

Reference: ``optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=0.0001)``
(/root/reference/part1/main.py:124-125).  Same constructor and ``zero_grad`` / ``step`` /
``state_dict`` surface; on CUDA ``step`` is ONE kernel over all 9.2M parameters
(csrc/sgd.cu) instead of a foreach loop over 34 tensors, followed by the bf16 filter repack.
``zero_grad`` is a no-op there: the backward kernels overwrite every gradient.
"""
from __future__ import annotations

import torch


class SGD(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 0.1, momentum: float = 0.0, weight_decay: float = 0.0):
        super().__init__(params, dict(lr=lr, momentum=momentum, weight_decay=weight_decay, dampening=0,
                                      nesterov=False))
        self._engine = None
        self._fallback = None

    def bind(self, engine) -> "SGD":
        """Attach the engine whose arenas hold these parameters."""
        self._engine = engine
        return self

    @property
    def hyper(self):
        g = self.param_groups[0]
        return g["lr"], g["momentum"], g["weight_decay"]

    def zero_grad(self, set_to_none: bool = True) -> None:
        if self._engine is not None:
            return
        super().zero_grad(set_to_none=set_to_none)

    @torch.no_grad()
    def step(self, closure=None):
        if self._engine is None:
            p0 = self.param_groups[0]["params"][0]
            eng = None
            if p0.is_cuda:
                # parameters re-homed by a VGGEngine share one storage: find it via the registry
                from .engine import engine_of_parameter

                eng = engine_of_parameter(p0)
            if eng is not None:
                self._engine = eng
        if self._engine is not None:
            lr, mu, wd = self.hyper
            self._engine.sgd(lr, mu, wd)   # also refreshes the bf16 shadow the conv kernels read
            return None
        lr, mu, wd = self.hyper
        if self._fallback is None:
            self._fallback = torch.optim.SGD(self.param_groups[0]["params"], lr=lr, momentum=mu, weight_decay=wd)
        for g in self._fallback.param_groups:   # follow later changes of our own param group (LR schedules)
            g["lr"], g["momentum"], g["weight_decay"] = lr, mu, wd
        return self._fallback.step(closure)

    def state_dict(self):
        sd = super().state_dict()
        if self._engine is not None:
            a = self._engine.arena
            sd["state"] = {i: {"momentum_buffer": a._view(a.momentum, p, o).clone(memory_format=torch.contiguous_format)}
                           for i, (p, o) in enumerate(zip(a.params, a.offsets))}
        elif self._fallback is not None:
            sd["state"] = self._fallback.state_dict()["state"]
        return sd

    def load_state_dict(self, sd):
        state = sd.get("state", {})
        if self._engine is not None and state:
            a = self._engine.arena
            for i, (p, o) in enumerate(zip(a.params, a.offsets)):
                buf = state.get(i, state.get(str(i)))
                if buf is not None:
                    a._view(a.momentum, p, o).copy_(buf["momentum_buffer"])
        elif state:
            lr, mu, wd = self.hyper
            self._fallback = torch.optim.SGD(self.param_groups[0]["params"], lr=lr, momentum=mu, weight_decay=wd)
            self._fallback.load_state_dict({"state": state, "param_groups": self._fallback.state_dict()["param_groups"]})


class CrossEntropyLoss(torch.nn.CrossEntropyLoss):
    """Drop-in for the reference's criterion (part1/main.py:119).  Kept as a module so the
    reference's call shape ``criterion(output, target)`` works; inside Trainer.step the loss is
    produced by the fused classifier kernel instead (csrc/fc_ce.cu)."""
