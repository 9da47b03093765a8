"""Flat-arena optimizers: one kernel per step over the whole d-length parameter vector.

`FlatOptimizer` keeps `torch.optim`-compatible hyper-parameters (`param_groups[0]['lr']` etc.,
so `adjust_learning_rate` and LR schedulers keep working) but its state lives in flat fp32
buffers next to the parameter arena. On CPU tensors the same arithmetic runs in torch.
Reference: /root/reference/pytorch_impl/libs/garfieldpp/tools.py:107-123 (name → torch.optim
class) and server.py:277-287 (flat gradient → per-param .grad → optimizer.step on the host).
"""

from __future__ import annotations

import ctypes

import torch

from . import _lib

__all__ = ["FlatOptimizer", "OPTIMIZERS"]

OPTIMIZERS = ("sgd", "adam", "adamw", "rmsprop", "adagrad")

_DEFAULTS = {
    "sgd": dict(lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False),
    "adam": dict(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0),
    "adamw": dict(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2),
    "rmsprop": dict(lr=1e-2, alpha=0.99, eps=1e-8, weight_decay=0.0, momentum=0.0),
    "adagrad": dict(lr=1e-2, lr_decay=0.0, weight_decay=0.0, eps=1e-10),
}


class FlatOptimizer(torch.optim.Optimizer):
    """Optimizer over one flat parameter tensor. `step(flat_grad)` applies one update in place.

    Subclasses `torch.optim.Optimizer` so LR schedulers, `adjust_learning_rate` and
    `param_groups` edits behave as with the reference's torch optimizers.
    """

    def __init__(self, flat_param, name, *args, **kwargs):
        if name not in _DEFAULTS:
            raise KeyError(name)
        hyper = dict(_DEFAULTS[name])
        if args:  # positional lr, like torch.optim constructors
            hyper["lr"] = args[0]
        for key, value in kwargs.items():
            if key not in hyper:
                raise TypeError("unexpected optimizer argument %r for %s" % (key, name))
            hyper[key] = value
        self.name = name
        self.param = flat_param
        self.steps = 0
        super().__init__([flat_param], hyper)

    # -- state helpers ---------------------------------------------------------------------
    def _buf(self, key, fill=0.0):
        slot = self.state[self.param]
        if key not in slot:
            slot[key] = torch.full_like(self.param, fill)
        return slot[key]

    @property
    def hyper(self):
        return self.param_groups[0]

    def zero_grad(self, set_to_none=True):
        return None  # gradients are never stored on the flat parameter

    def state_dict(self):
        hyper = {k: v for k, v in self.hyper.items() if k != "params"}
        return {"name": self.name, "hyper": hyper, "steps": self.steps,
                "state": {k: v.detach().clone() for k, v in self.state[self.param].items()}}

    def load_state_dict(self, sd):
        if sd["name"] != self.name:
            raise ValueError("optimizer kind mismatch: %s vs %s" % (sd["name"], self.name))
        self.hyper.update(sd["hyper"])
        self.steps = int(sd["steps"])
        slot = self.state[self.param]
        slot.clear()
        for k, v in sd["state"].items():
            slot[k] = v.to(self.param.device).clone()

    def sgd_fusion(self, mirrors=(), multicast=None):
        """`ops.gar.SgdFusion` describing this optimizer's next step (SGD only). `multicast`: NVLS multicast address of the
        parameter slice — publication to every GPU by one multimem.st instead of one store per mirror."""
        from .gar import SgdFusion
        if self.name != "sgd":
            raise ValueError("only SGD can be fused into the aggregation epilogue")
        h = self.hyper
        mom = self._buf("momentum_buffer") if h["momentum"] != 0 else None
        return SgdFusion(self.param, mom, h["lr"], h["momentum"], h["dampening"], h["weight_decay"], h["nesterov"],
                         first_step=(self.steps == 0), mirrors=() if multicast else mirrors, multicast=multicast)

    def note_fused_step(self):
        self.steps += 1

    def rebind(self, flat_param):
        """Continue optimising another flat tensor of the same shape (double-buffered arenas):
        hyper-parameters, step count and state buffers are kept."""
        if flat_param.shape != self.param.shape:
            raise ValueError("rebind needs a tensor of the same shape")
        slot = self.state.pop(self.param, {})
        self.param = flat_param
        self.param_groups[0]["params"] = [flat_param]
        self.state[flat_param] = slot
        return self

    # -- the update ------------------------------------------------------------------------
    @torch.no_grad()
    def step(self, grad=None, mirrors=()):
        p, h = self.param, self.hyper
        if grad is None or callable(grad):  # torch-style call: use the gradient attached to the flat parameter
            grad = p.grad
            if grad is None:
                return None
        grad = grad.to(device=p.device, dtype=p.dtype).reshape(-1)
        self.steps += 1
        use_kernel = p.is_cuda and p.dtype == torch.float32 and _lib.available()
        if use_kernel:
            self._step_cuda(grad.contiguous(), mirrors)
        else:
            self._step_torch(grad)
            for mirror in mirrors:
                mirror.copy_(p)

    def _step_cuda(self, g, mirrors):
        lib, p, h = _lib.require(), self.param, self.hyper
        d, st = p.numel(), _lib.stream_ptr(p.device)
        with torch.cuda.device(p.device):
            if self.name == "sgd":
                mom = self._buf("momentum_buffer") if h["momentum"] != 0 else None
                table = (ctypes.c_void_p * max(1, len(mirrors)))(*[m.data_ptr() if isinstance(m, torch.Tensor) else int(m) for m in mirrors])
                _lib.check(lib.gb_sgd_step(p.data_ptr(), g.data_ptr(), mom.data_ptr() if mom is not None else None, d,
                                           h["lr"], h["momentum"], h["dampening"], h["weight_decay"], int(h["nesterov"]),
                                           int(self.steps == 1), table, len(mirrors), st))
            elif self.name in ("adam", "adamw"):
                _lib.check(lib.gb_adam_step(p.data_ptr(), g.data_ptr(), self._buf("exp_avg").data_ptr(), self._buf("exp_avg_sq").data_ptr(),
                                            d, h["lr"], h["betas"][0], h["betas"][1], h["eps"], h["weight_decay"], self.steps,
                                            int(self.name == "adamw"), st))
            elif self.name == "rmsprop":
                mom = self._buf("momentum_buffer") if h["momentum"] != 0 else None
                _lib.check(lib.gb_rmsprop_step(p.data_ptr(), g.data_ptr(), self._buf("square_avg").data_ptr(),
                                               mom.data_ptr() if mom is not None else None, d, h["lr"], h["alpha"], h["eps"],
                                               h["weight_decay"], h["momentum"], st))
            else:
                _lib.check(lib.gb_adagrad_step(p.data_ptr(), g.data_ptr(), self._buf("sum").data_ptr(), d, h["lr"], h["lr_decay"],
                                               h["eps"], h["weight_decay"], self.steps, st))
        _lib.count()
        if self.name != "sgd":
            for mirror in mirrors:
                mirror.copy_(p)

    def _step_torch(self, g):
        p, h = self.param, self.hyper
        if self.name == "sgd":
            if h["weight_decay"] != 0:
                g = g.add(p, alpha=h["weight_decay"])
            if h["momentum"] != 0:
                buf = self._buf("momentum_buffer")
                if self.steps == 1:
                    buf.copy_(g)
                else:
                    buf.mul_(h["momentum"]).add_(g, alpha=1 - h["dampening"])
                g = g.add(buf, alpha=h["momentum"]) if h["nesterov"] else buf
            p.add_(g, alpha=-h["lr"])
        elif self.name in ("adam", "adamw"):
            b1, b2 = h["betas"]
            if self.name == "adamw":
                p.mul_(1 - h["lr"] * h["weight_decay"])
            elif h["weight_decay"] != 0:
                g = g.add(p, alpha=h["weight_decay"])
            m, v = self._buf("exp_avg"), self._buf("exp_avg_sq")
            m.mul_(b1).add_(g, alpha=1 - b1)
            v.mul_(b2).addcmul_(g, g, value=1 - b2)
            bc1, bc2 = 1 - b1 ** self.steps, 1 - b2 ** self.steps
            denom = (v.sqrt() / (bc2 ** 0.5)).add_(h["eps"])
            p.addcdiv_(m, denom, value=-h["lr"] / bc1)
        elif self.name == "rmsprop":
            if h["weight_decay"] != 0:
                g = g.add(p, alpha=h["weight_decay"])
            sq = self._buf("square_avg")
            sq.mul_(h["alpha"]).addcmul_(g, g, value=1 - h["alpha"])
            avg = sq.sqrt().add_(h["eps"])
            if h["momentum"] != 0:
                buf = self._buf("momentum_buffer")
                buf.mul_(h["momentum"]).addcdiv_(g, avg)
                p.add_(buf, alpha=-h["lr"])
            else:
                p.addcdiv_(g, avg, value=-h["lr"])
        else:
            if h["weight_decay"] != 0:
                g = g.add(p, alpha=h["weight_decay"])
            clr = h["lr"] / (1 + (self.steps - 1) * h["lr_decay"])
            s = self._buf("sum")
            s.addcmul_(g, g, value=1.0)
            p.addcdiv_(g, s.sqrt().add_(h["eps"]), value=-clr)
