"""Flat device arenas: a model whose parameters are views of ONE contiguous fp32 vector.

This is the framework's memory layout: parameters, per-worker gradients and
optimizer state are flat d-length device buffers in `parameters()` order with no padding — the
same order the reference uses when it flattens (`garfieldpp/worker.py:93-94`,
`server.py:200,283-297`), so flat vectors are interchangeable with the reference's and the
`state_dict` ↔ flat mapping is a bijection (checkpoints, §5.4). Because parameters are views,
"flatten" and "unflatten" never copy: `update_model`/`write_model` operate on the vector and
the module sees the result immediately; gradients land directly in a row of the [n, d]
gradient matrix the aggregation kernels read.
"""

from __future__ import annotations

import torch

__all__ = ["FlatModel", "flat_size", "GradMatrix"]


def flat_size(module):
    return sum(p.numel() for p in module.parameters())


class FlatModel:
    """Wrap `module` so all its parameters alias `self.flat` (float32, on `device`)."""

    def __init__(self, module, device=None, flat=None, channels_last=False):
        """`channels_last=True` stores every 4-D parameter in OHWI (NHWC) physical order inside the flat
        vector — cuDNN's fast layout on sm_100 (2× faster ResNet-50 steps here). Logical shapes,
        `state_dict`s and checkpoints are unchanged; only the element order INSIDE the flat vector
        differs from the reference's, which no aggregation rule can observe (they are invariant
        under a fixed permutation of coordinates). `to_reference_order` / `from_reference_order`
        convert flat vectors at API boundaries."""
        if device is not None:
            module = module.to(device)
        self.module = module
        self.channels_last = bool(channels_last)
        params = list(module.parameters())
        if not params:
            raise ValueError("model has no parameters")
        self.device = params[0].device
        self.slices = []
        offset = 0
        for p in params:
            self.slices.append((offset, p.numel(), tuple(p.shape)))
            offset += p.numel()
        self.numel = offset
        if flat is None:
            flat = torch.empty(self.numel, dtype=torch.float32, device=self.device)
            self._adopt(flat, copy_from_params=True)
        else:
            if flat.numel() != self.numel or flat.dtype != torch.float32:
                raise ValueError("flat buffer must be float32 with %d elements" % self.numel)
            self._adopt(flat, copy_from_params=False)

    def _view(self, vec, off, n, shape):
        chunk = vec[off:off + n]
        if self.channels_last and len(shape) == 4:
            o, i, h, w = shape
            return chunk.view(o, h, w, i).permute(0, 3, 1, 2)  # logical OIHW, physical OHWI
        return chunk.view(shape)

    def _adopt(self, flat, copy_from_params):
        with torch.no_grad():
            for p, (off, n, shape) in zip(self.module.parameters(), self.slices):
                view = self._view(flat, off, n, shape)
                if copy_from_params:
                    view.copy_(p.data)
                p.data = view
        self.flat = flat

    # -- module passthrough ---------------------------------------------------------------
    def __call__(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def parameters(self):
        return self.module.parameters()

    def train(self, mode=True):
        self.module.train(mode)
        return self

    def eval(self):
        self.module.eval()
        return self

    def state_dict(self):
        return self.module.state_dict()

    def load_state_dict(self, sd, strict=True):
        # copy INTO the views so the aliasing with `flat` survives
        own = self.module.state_dict()
        missing = [k for k in own if k not in sd]
        unexpected = [k for k in sd if k not in own]
        if strict and (missing or unexpected):
            raise KeyError("state_dict mismatch: missing %s, unexpected %s" % (missing, unexpected))
        with torch.no_grad():
            for key, tensor in own.items():
                if key in sd:
                    tensor.copy_(sd[key])

    # -- flat access -----------------------------------------------------------------------
    def views_of(self, vec):
        """Per-parameter views (parameter shapes) into another flat vector of length d."""
        return [self._view(vec, off, n, shape) for off, n, shape in self.slices]

    def to_reference_order(self, vec):
        """Copy of a flat vector in the reference's element order (each parameter row-major OIHW)."""
        if not self.channels_last:
            return vec.clone()
        return torch.cat([v.contiguous().reshape(-1) for v in self.views_of(vec)])

    def from_reference_order(self, vec):
        """Inverse of `to_reference_order`: a flat vector in this model's internal order."""
        if not self.channels_last:
            return vec.clone()
        out = torch.empty_like(vec)
        for (off, n, shape), view in zip(self.slices, self.views_of(out)):
            view.copy_(vec[off:off + n].view(shape))
        return out

    def bind_grads(self, vec):
        """Make every `param.grad` a view of `vec` (autograd then accumulates in place)."""
        for p, view in zip(self.module.parameters(), self.views_of(vec)):
            p.grad = view

    def unbind_grads(self):
        for p in self.module.parameters():
            p.grad = None

    def write(self, vec):
        """Overwrite all parameters from a flat vector (any device)."""
        with torch.no_grad():
            self.flat.copy_(vec.reshape(-1))

    def buffers_state(self):
        """Non-parameter state (BN running statistics …) as a name → tensor dict."""
        return {k: v for k, v in self.module.named_buffers()}


class GradMatrix:
    """[n, d] float32 matrix whose rows are the workers' flat gradients.

    Rows are 256-byte aligned (d is padded to a multiple of 64 floats internally) so every row is
    a valid TMA bulk-copy source for the aggregation kernels.
    """

    def __init__(self, n, d, device, storage=None):
        self.n, self.d = n, d
        self.pitch = (d + 63) // 64 * 64
        if storage is None:
            storage = torch.zeros(n * self.pitch, dtype=torch.float32, device=device)
        self.storage = storage
        self.rows = [storage[i * self.pitch:i * self.pitch + d] for i in range(n)]

    def row(self, i):
        return self.rows[i]

    def zero_(self):
        self.storage.zero_()
