"""Import harness for the UNMODIFIED reference Python under /root/reference.

Test infrastructure only (used by tests/golden/gen_*.py in the authoring container to
produce the committed golden fixtures, and by `-m "not gpu"` tests that cross-check the
oracle restatement when /root/reference is present).  Never imported by the product.

The reference (Isaac Lab 2.1.0 + rsl_rl 3.0.0) imports Omniverse/PhysX/Warp/trimesh/
gymnasium/tensordict at module scope.  None of these are needed for the arithmetic of the
hot path, so they are replaced by inert stub modules; `tensordict.TensorDict` gets a small
structural stand-in (no arithmetic).  SURVEY.md §7 step 0 / §8(c).
"""
from __future__ import annotations

import importlib.abc
import importlib.machinery
import os
import sys
import types

import torch

def _default_root():
    """/root/reference where it exists (authoring container); else the unmodified copy that baseline/install_ref.py put
    under baseline/_ref/ (the GPU box)."""
    if os.path.isdir("/root/reference/source/isaaclab"):
        return "/root/reference"
    here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    return os.path.join(here, "baseline", "_ref")


REFERENCE_ROOT = os.environ.get("B200LAB_REFERENCE_ROOT") or _default_root()

_STUB_ROOTS = (
    "omni", "carb", "isaacsim", "pxr", "prettytable", "warp", "trimesh", "gymnasium", "matplotlib",
    "flatdict", "hidapi", "hid", "mpl_toolkits", "h5py", "onnx", "pink", "pinocchio", "pyglet", "junitparser",
    "dex_retargeting", "Semantics", "usdrt", "nvidia.srl", "tensorboard", "torch.utils.tensorboard",
)


class _StubMeta(type):
    def __getattr__(cls, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _make_stub(f"{cls.__name__}.{name}")

    def __call__(cls, *a, **k):  # instantiating a stub yields an inert object
        return super().__call__()

    def __iter__(cls):
        return iter(())

    def __or__(cls, other):
        return cls

    def __ror__(cls, other):
        return cls


def _make_stub(name):
    return _StubMeta(name.split(".")[-1], (), {"__init__": lambda self, *a, **k: None,
                                               "__getattr__": lambda self, n: _make_stub(n)(),
                                               "__call__": lambda self, *a, **k: None,
                                               "__module__": "stub"})


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__") and name not in ("__path__",):
            raise AttributeError(name)
        full = f"{self.__name__}.{name}"
        if full in sys.modules:
            return sys.modules[full]
        stub = _make_stub(full)
        setattr(self, name, stub)
        return stub


class _StubFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        for root in _STUB_ROOTS:
            if fullname == root or fullname.startswith(root + "."):
                return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        mod = _StubModule(spec.name)
        mod.__path__ = []
        return mod

    def exec_module(self, module):
        pass


class TensorDict:
    """Structural stand-in for tensordict.TensorDict (what rsl_rl 3.0.0 touches, nothing more)."""

    def __init__(self, source=None, batch_size=None, device=None):
        self._d = dict(source or {})
        self.batch_size = torch.Size(batch_size if batch_size is not None else [])
        self.device = device

    # mapping surface
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

    @property
    def shape(self):
        return self.batch_size

    def __getitem__(self, idx):
        if isinstance(idx, str):
            return self._d[idx]
        out = {k: v[idx] for k, v in self._d.items()}
        any_v = next(iter(out.values()))
        nb = len(self.batch_size)
        ref = next(iter(self._d.values()))
        bs = any_v.shape[: any_v.dim() - (ref.dim() - nb)]
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
        return TensorDict({k: v.to(device) for k, v in self._d.items()}, batch_size=self.batch_size, device=device)

    def clone(self):
        return TensorDict({k: v.clone() for k, v in self._d.items()}, batch_size=self.batch_size, device=self.device)


_installed = False


def install():
    """Install stubs + sys.path entries so `import isaaclab`, `import rsl_rl` resolve to the reference."""
    global _installed
    if _installed:
        return
    if not os.path.isdir(REFERENCE_ROOT):
        raise FileNotFoundError(f"reference tree not found at {REFERENCE_ROOT}")
    sys.meta_path.insert(0, _StubFinder())
    td = types.ModuleType("tensordict")
    td.TensorDict = TensorDict
    sys.modules["tensordict"] = td
    # toml: the reference reads its extension.toml at import time -> real parser (tomllib)
    import tomllib
    toml = types.ModuleType("toml")
    toml.load = lambda path: tomllib.load(open(path, "rb"))
    sys.modules["toml"] = toml
    # gymnasium needs real classes for `class X(gym.Env)` and gym.spaces.* calls at import time
    import importlib
    gym = importlib.import_module("gymnasium")
    gym.Env = type("Env", (), {})
    gym.Wrapper = type("Wrapper", (), {})
    for sub in ("isaaclab", "isaaclab_tasks", "isaaclab_assets", "isaaclab_rl", "rsl_rl"):
        p = os.path.join(REFERENCE_ROOT, "source", sub)
        if p not in sys.path:
            sys.path.insert(0, p)
    # isaaclab_tasks/__init__.py walks and imports every task package (many need Kit): bypass it with
    # a bare package object so that only the sub-modules we ask for are imported.
    tasks = types.ModuleType("isaaclab_tasks")
    tasks.__path__ = [os.path.join(REFERENCE_ROOT, "source", "isaaclab_tasks", "isaaclab_tasks")]
    sys.modules["isaaclab_tasks"] = tasks
    _installed = True


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "source", "isaaclab"))
