"""The drop-in boundary without a GPU: libb200lab.so loads, exports every symbol include/b200lab.h declares, the ctypes
mirror of every struct has the C compiler's size, and host-side argument checks fail loudly (no compute calls here)."""
import ctypes
import os
import re

import pytest

from isaaclab_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def L():
    if not _lib.available():
        from isaaclab_b200 import build
        build.build()
    return _lib.lib()


def test_every_declared_symbol_is_exported(L):
    names = _lib.declared_symbols()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, f"not exported: {missing}"


def test_every_declared_symbol_has_a_ctypes_signature():
    missing = [n for n in _lib.declared_symbols() if n not in _lib._SIGS]
    assert not missing, f"no ctypes signature: {missing}"


def test_struct_sizes_match_the_c_compiler(L):
    for name, cls in _lib.structs.items():
        c_size = L.b200lab_sizeof(name.encode())
        assert c_size == ctypes.sizeof(cls), f"{name}: C sizeof={c_size}, ctypes={ctypes.sizeof(cls)}"


def test_header_has_no_torch_types():
    src = _lib._strip_comments(open(os.path.join(ROOT, "include", "b200lab.h")).read())
    assert not re.search(r"(at::|torch::|c10::|\bTensor\b|#include\s*<(torch|ATen|c10))", src)
    assert 'extern "C"' in src
    assert set(re.findall(r"#include\s*<([^>]+)>", src)) <= {"stdint.h", "stddef.h"}


def test_version_and_error_string(L):
    assert L.b200lab_version() >= 1
    assert isinstance(L.b200lab_last_error(), bytes)


def test_null_arguments_are_rejected_on_the_host(L):
    assert L.b200lab_process_action(None, None, None) != 0
    assert b"null" in L.b200lab_last_error().lower()
    assert L.b200lab_gae(None, None, None, None, None, None, 0, 0, 0.99, 0.95, 0, None, 0, None) != 0


def test_ptr_rejects_host_tensors():
    import torch
    with pytest.raises((ValueError, RuntimeError, TypeError)):
        _lib.ptr(torch.zeros(4))


def test_no_product_module_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under isaaclab_b200/ may import it."""
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "isaaclab_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", src, re.M):
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad


def test_integration_md_binding_stub_is_generated_from_the_header_and_executes(L):
    """INTEGRATION.md section 1 is the output of tools/gen_integration.py (no drift), it executes against the built
    library, and its signatures agree with the product's own binding table (two independent derivations of the header)."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("gen_integration", os.path.join(ROOT, "tools", "gen_integration.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    assert gen.render(doc) == doc, "INTEGRATION.md is stale: run python tools/gen_integration.py"
    block = doc[doc.index(gen.BEGIN):doc.index(gen.END)]
    code = block[block.index("```python") + len("```python"):block.rindex("```")]
    code = code.replace('C.CDLL("libb200lab.so")', f'C.CDLL({_lib.LIB_PATH!r})')
    ns = {}
    exec(compile(code, "INTEGRATION.md", "exec"), ns)
    sigs = ns["_SIGS"]
    assert set(sigs) == set(_lib.declared_symbols()) == set(_lib._SIGS)
    for name, (res, args) in sigs.items():
        want_res, want_args = _lib._SIGS[name]
        assert res is want_res, f"{name}: restype {res} vs {want_res}"
        assert len(args) == len(want_args), f"{name}: {len(args)} arguments documented, {len(want_args)} bound"
        for i, (a, b) in enumerate(zip(args, want_args)):
            assert a is b, f"{name} argument {i}: documented {a}, bound {b}"
    assert ns["_L"].b200lab_version() >= 1 and ns["ptr"](None) is None
