"""`TensorDict` as used by the rsl_rl 3.0.0 interfaces (RSL/env/vec_env.py:49, RSL/storage/rollout_storage.py:48-52).

The real `tensordict` package is used when installed; otherwise this structural container (a dict of tensors
sharing leading batch dims -- no arithmetic) provides the surface the runner/storage/policy touch.
"""
from __future__ import annotations

import torch

try:  # pragma: no cover - depends on the image
    from tensordict import TensorDict  # type: ignore
except ImportError:

    class TensorDict:  # noqa: D101
        def __init__(self, source=None, batch_size=None, device=None):
            self._d = dict(source or {})
            self.batch_size = torch.Size(batch_size if batch_size is not None else [])
            self.device = device

        def keys(self):
            return self._d.keys()

        def items(self):
            return self._d.items()

        def values(self):
            return self._d.values()

        def __contains__(self, k):
            return k in self._d

        def __iter__(self):
            return iter(self._d)

        def __len__(self):
            return len(self._d)

        @property
        def shape(self):
            return self.batch_size

        def __getitem__(self, idx):
            if isinstance(idx, str):
                return self._d[idx]
            nb = len(self.batch_size)
            out, bs = {}, None
            for k, v in self._d.items():
                o = v[idx]
                out[k] = o
                if bs is None:
                    bs = o.shape[: max(0, o.dim() - (v.dim() - nb))]
            return TensorDict(out, batch_size=bs, device=self.device)

        def __setitem__(self, idx, value):
            if isinstance(idx, str):
                self._d[idx] = value
                return
            for k, v in self._d.items():
                v[idx] = value[k]

        def copy_(self, other):
            for k, v in self._d.items():
                v.copy_(other[k])
            return self

        def flatten(self, a, b):
            out = {k: v.flatten(a, b) for k, v in self._d.items()}
            bs = next(iter(out.values())).shape[: len(self.batch_size) - (b - a)]
            return TensorDict(out, batch_size=bs, device=self.device)

        def to(self, device):
            if all(v.device == torch.device(device) for v in self._d.values()):
                return self
            return TensorDict({k: v.to(device) for k, v in self._d.items()}, batch_size=self.batch_size, device=device)

        def clone(self):
            return TensorDict({k: v.clone() for k, v in self._d.items()}, batch_size=self.batch_size,
                              device=self.device)
