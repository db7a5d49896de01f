"""ctypes binding of libb200lab.so (the C ABI declared in include/b200lab.h).

The ctypes `Structure` mirrors and the enum constants are GENERATED from the header at import time, so the
Python side cannot drift from the C side; `b200lab_sizeof` cross-checks every struct size.

There is no CPU fallback: if the library is missing, `lib()` raises; if a call fails, `check()` raises
`RuntimeError` with `b200lab_last_error()`.
"""
from __future__ import annotations

import ctypes as C
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(HERE, "..", "include", "b200lab.h")
LIB_PATH = os.path.join(HERE, "libb200lab.so")

_CTYPES = {
    "int32_t": C.c_int32, "int64_t": C.c_int64, "uint64_t": C.c_uint64, "uint8_t": C.c_uint8, "int8_t": C.c_int8,
    "float": C.c_float, "double": C.c_double, "int": C.c_int, "void": None,
}

consts: dict[str, int] = {}
structs: dict[str, type] = {}


def _strip_comments(src: str) -> str:
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.sub(r"//[^\n]*", "", src)


def _parse_header():
    src = _strip_comments(open(HEADER).read())
    # #define NAME int
    for m in re.finditer(r"#define\s+(B200LAB_\w+)\s+\(?(-?\d+)\)?\s*$", src, flags=re.M):
        consts[m.group(1)] = int(m.group(2))
    # enums
    for m in re.finditer(r"enum\s+\w+\s*\{(.*?)\}", src, flags=re.S):
        nxt = 0
        for item in m.group(1).split(","):
            item = item.strip()
            if not item:
                continue
            if "=" in item:
                k, v = item.split("=")
                nxt = int(v.strip(), 0)
                consts[k.strip()] = nxt
            else:
                consts[item] = nxt
            nxt += 1
    # structs (in order; may reference earlier structs)
    for m in re.finditer(r"typedef\s+struct\s+(\w+)\s*\{(.*?)\}\s*\1\s*;", src, flags=re.S):
        name, body = m.group(1), m.group(2)
        fields = []
        for decl in body.split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            decl = decl.replace("const ", "")
            mm = re.match(r"(\w+)\s*(.*)", decl)
            base, rest = mm.group(1), mm.group(2)
            for var in rest.split(","):
                var = var.strip()
                ptr = var.count("*")
                var = var.replace("*", "").strip()
                dims = [int(consts.get(d, d)) if not d.isdigit() else int(d) for d in re.findall(r"\[(\w+)\]", var)]
                vname = re.match(r"\w+", var).group(0)
                if ptr:
                    ct = C.c_void_p
                elif base in structs:
                    ct = structs[base]
                else:
                    ct = _CTYPES[base]
                for d in reversed(dims):
                    ct = ct * d
                fields.append((vname, ct))
        structs[name] = type(name, (C.Structure,), {"_fields_": fields})


_parse_header()

_lib = None

_SIGS = {
    "b200lab_last_error": (C.c_char_p, []),
    "b200lab_version": (C.c_int, []),
    "b200lab_launch_count": (C.c_int64, []),
    "b200lab_sizeof": (C.c_int64, [C.c_char_p]),
    "b200lab_env_create": (C.c_void_p, [C.c_void_p]),
    "b200lab_env_update": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200lab_env_destroy": (None, [C.c_void_p]),
    "b200lab_process_action": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200lab_substep_pre": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200lab_substep_post": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200lab_decimation_fused": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "b200lab_height_scan_prefetch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200lab_post_step": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "b200lab_post_reset_write": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200lab_step_tail": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "b200lab_compute_obs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "b200lab_log_fold": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200lab_compute_obs_range": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_void_p]),
    "b200lab_step_tail_begin": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200lab_step_tail_end": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200lab_lstm_cell_fwd": (C.c_int, [C.c_void_p] * 6 + [C.c_int64, C.c_int32, C.c_void_p]),
    "b200lab_lstm_cell_bwd": (C.c_int, [C.c_void_p] * 7 + [C.c_int64, C.c_int32, C.c_void_p]),
    "b200lab_gru_cell_fwd": (C.c_int, [C.c_void_p] * 5 + [C.c_int64, C.c_int32, C.c_void_p]),
    "b200lab_gru_cell_bwd": (C.c_int, [C.c_void_p] * 8 + [C.c_int64, C.c_int32, C.c_void_p]),
    "b200lab_mesh_create": (C.c_void_p, [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_float]),
    "b200lab_mesh_device_ptr": (C.c_void_p, [C.c_void_p]),
    "b200lab_mesh_bytes": (C.c_int64, [C.c_void_p]),
    "b200lab_mesh_destroy": (None, [C.c_void_p]),
    "b200lab_raycast_mesh": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200lab_gae_scratch_bytes": (C.c_int64, [C.c_int32, C.c_int32]),
    "b200lab_gae": (C.c_int, [C.c_void_p] * 6 + [C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_int32,
                              C.c_void_p, C.c_void_p]),
    "b200lab_gae_static": (C.c_int, [C.c_void_p] * 6 + [C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_int32, C.c_void_p]),
    "b200lab_add_transition": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200lab_empirical_norm_update": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int64] + [C.c_void_p] * 4),
    "b200lab_empirical_norm_partial": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]),
    "b200lab_empirical_norm_merge": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64] + [C.c_void_p] * 4),
    "b200lab_empirical_norm_apply": (C.c_int, [C.c_void_p] * 3 + [C.c_float, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    "b200lab_episode_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32] + [C.c_void_p] * 4
                              + [C.c_int32, C.c_void_p, C.c_void_p]),
    "b200lab_gather_minibatch": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200lab_linear_fwd": (C.c_int, [C.c_void_p] * 4 + [C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                     C.c_void_p]),
    "b200lab_linear_bwd_workspace_bytes": (C.c_int64, [C.c_int64, C.c_int32, C.c_int32]),
    "b200lab_linear_bwd": (C.c_int, [C.c_void_p] * 8 + [C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                     C.c_void_p, C.c_void_p]),
    "b200lab_linear_bwd_chain": (C.c_int, [C.c_void_p] * 9 + [C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                           C.c_void_p, C.c_void_p]),
    "b200lab_sample_actions": (C.c_int, [C.c_void_p] * 6 + [C.c_int64, C.c_int32, C.c_void_p]),
    "b200lab_ppo_loss": (C.c_int, [C.c_void_p, C.c_void_p]),
    "b200lab_adapt_lr": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p]),
    "b200lab_grad_sumsq": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "b200lab_adam_clip_step": (C.c_int, [C.c_void_p] * 4 + [C.c_int64, C.c_void_p, C.c_void_p, C.c_float, C.c_float,
                                         C.c_float, C.c_float, C.c_float, C.c_int64, C.c_void_p]),
}


def declared_symbols() -> list[str]:
    """Every function name declared in include/b200lab.h."""
    src = _strip_comments(open(HEADER).read())
    return sorted(set(re.findall(r"\b(b200lab_\w+)\s*\(", src)))


def lib() -> C.CDLL:
    """Load libb200lab.so (built in-tree by isaaclab_b200.build).  Raises if absent -- there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m isaaclab_b200.build` "
            "(isaaclab_b200 has no CPU / PyTorch fallback for its CUDA path)")
    l = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGS.items():
        fn = getattr(l, name)
        fn.restype = res
        fn.argtypes = args
    for name, st in structs.items():
        n = l.b200lab_sizeof(name.encode())
        if n != C.sizeof(st):
            raise RuntimeError(f"ABI mismatch for {name}: C sizeof={n}, ctypes sizeof={C.sizeof(st)}")
    _lib = l
    return l


def available() -> bool:
    """True when the built library file is present (does not load it)."""
    return os.path.exists(LIB_PATH)


def check(rc: int):
    if rc != 0:
        raise RuntimeError(f"libb200lab error {rc}: {lib().b200lab_last_error().decode()}")


def ptr(t) -> int | None:
    """Raw device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("libb200lab operates on CUDA tensors only (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("libb200lab requires contiguous tensors")
    return t.data_ptr()


def current_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
