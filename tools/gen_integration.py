"""Generates the ctypes binding stub of INTEGRATION.md section 1 from the prototypes in include/b200lab.h, so the
documented signatures cannot drift from the header (tests/test_cabi.py regenerates and executes it).

    python tools/gen_integration.py           rewrites the block between the BEGIN/END GENERATED markers in INTEGRATION.md
"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200lab.h")
DOC = os.path.join(ROOT, "INTEGRATION.md")
BEGIN, END = "<!-- BEGIN GENERATED: binding stub (tools/gen_integration.py) -->", "<!-- END GENERATED -->"

_SCALARS = {"int32_t": "C.c_int32", "int64_t": "C.c_int64", "uint64_t": "C.c_uint64", "float": "C.c_float",
            "double": "C.c_double", "int": "C.c_int", "uint8_t": "C.c_uint8"}


def _ctype(decl: str, is_return=False) -> str:
    d = decl.strip()
    if "*" in d:
        base = d.replace("const", "").replace("*", " ").split()[0]
        return "C.c_char_p" if base == "char" else "C.c_void_p"
    base = d.replace("const", "").split()[0]
    if base == "void":
        return "None"
    return _SCALARS[base]


def prototypes():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    src = re.sub(r"^\s*#[^\n]*$", "", src, flags=re.M)          # preprocessor lines
    src = src.replace('extern "C" {', "")
    src = re.sub(r"enum\s+\w+\s*\{.*?\}\s*;", "", src, flags=re.S)
    src = re.sub(r"typedef\s+struct\s+\w+\s*\{.*?\}\s*\w+\s*;", "", src, flags=re.S)
    out = []
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(b200lab_\w+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
        ret, name, args = m.group(1).strip(), m.group(2), " ".join(m.group(3).split())
        params = [] if args in ("", "void") else [a.strip() for a in args.split(",")]
        out.append((name, ret, params))
    return out


def stub() -> str:
    lines = ['import ctypes as C', '', '_L = C.CDLL("libb200lab.so")   # isaaclab_b200/libb200lab.so (python -m isaaclab_b200.build)', '',
             '# name: (restype, [argtypes])  <-  include/b200lab.h; every pointer is a raw DEVICE pointer unless the header says host',
             '_SIGS = {']
    for name, ret, params in prototypes():
        names = ", ".join(p.replace("const ", "").replace("*", "* ").split()[-1] for p in params)
        lines.append(f'    "{name}": ({_ctype(ret, True)}, [{", ".join(_ctype(p) for p in params)}]),   # ({names})')
    lines += ['}', 'for _name, (_res, _args) in _SIGS.items():', '    _fn = getattr(_L, _name)', '    _fn.restype, _fn.argtypes = _res, _args', '', '',
              'def check(rc):', '    if rc:', '        raise RuntimeError(_L.b200lab_last_error().decode())', '', '',
              'def ptr(t):   # contiguous CUDA tensor -> raw device pointer (None -> NULL); keep `t` alive until the stream consumed it',
              '    if t is None:', '        return None', '    assert t.is_cuda and t.is_contiguous()', '    return t.data_ptr()']
    return "\n".join(lines)


def render(doc: str) -> str:
    i0, i1 = doc.index(BEGIN), doc.index(END)
    return doc[:i0] + BEGIN + "\n```python\n" + stub() + "\n```\n" + doc[i1:]


if __name__ == "__main__":
    doc = open(DOC).read()
    new = render(doc)
    if "--check" in sys.argv:
        sys.exit(0 if new == doc else 1)
    open(DOC, "w").write(new)
    print(f"{len(prototypes())} prototypes written to INTEGRATION.md")
