"""ctypes binding of libdexstep.so (the C ABI declared in include/dexstep.h).

The struct mirrors, enum values and function list are *parsed from the header* at import, so the
binding cannot drift from the C layout; ``dex_sizeof_params()`` is checked against
``ctypes.sizeof`` when the library is loaded.

There is no fallback: if the shared library is missing or cannot be loaded, ``load()`` raises.
"""
from __future__ import annotations

import ctypes
import os
import re
from typing import Dict, List, Tuple

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
_REPO_ROOT = os.path.dirname(_PKG_DIR)
HEADER_PATH = os.path.join(_REPO_ROOT, "include", "dexstep.h")
LIB_PATH = os.path.join(_PKG_DIR, "libdexstep.so")

_CTYPES = {
    "int32_t": ctypes.c_int32, "uint32_t": ctypes.c_uint32, "int64_t": ctypes.c_int64,
    "uint64_t": ctypes.c_uint64, "uint8_t": ctypes.c_uint8, "float": ctypes.c_float,
    "double": ctypes.c_double, "int": ctypes.c_int, "void": None,
}


def _strip_comments(src: str) -> str:
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    return re.sub(r"//[^\n]*", " ", src)


def _parse_header(path: str):
    with open(path) as f:
        src = _strip_comments(f.read())
    consts: Dict[str, int] = {}
    for m in re.finditer(r"#define\s+(DEX_\w+)\s+(-?\d+)u?\s*$", src, flags=re.M):
        consts[m.group(1)] = int(m.group(2))
    for m in re.finditer(r"enum\s+\w+\s*\{(.*?)\}\s*;", src, flags=re.S):
        val = -1
        for item in m.group(1).split(","):
            item = item.strip()
            if not item:
                continue
            if "=" in item:
                name, expr = [x.strip() for x in item.split("=")]
                val = consts[expr] if expr in consts else int(expr)
            else:
                name, val = item, val + 1
            consts[name] = val
    structs: Dict[str, List[Tuple[str, object]]] = {}
    for m in re.finditer(r"typedef\s+struct\s+(\w+)\s*\{(.*?)\}\s*\1\s*;", src, flags=re.S):
        fields = []
        for decl in m.group(2).split(";"):
            decl = decl.strip()
            if not decl:
                continue
            dm = re.match(r"(const\s+)?(\w+)\s*(\*?)\s*(.*)$", decl, flags=re.S)
            base, is_ptr, rest = dm.group(2), bool(dm.group(3)), dm.group(4)
            for d in rest.split(","):
                d = d.strip()
                ptr = is_ptr or d.startswith("*")
                d = d.lstrip("* ")
                am = re.match(r"(\w+)\s*\[\s*(\w+)\s*\]$", d)
                if ptr:
                    fields.append((d, ctypes.c_void_p))
                elif am:
                    dim = consts[am.group(2)] if am.group(2) in consts else int(am.group(2))
                    fields.append((am.group(1), _CTYPES[base] * dim))
                else:
                    fields.append((d, _CTYPES[base]))
        structs[m.group(1)] = fields
    funcs = re.findall(r"^\s*(?:const\s+char\s*\*|int)\s+(dex_\w+)\s*\(", src, flags=re.M)
    return consts, structs, funcs


CONSTS, _STRUCT_FIELDS, FUNCTIONS = _parse_header(HEADER_PATH)
globals().update(CONSTS)  # DEX_* names become module attributes


def _make_struct(name):
    return type(name, (ctypes.Structure,), {"_fields_": _STRUCT_FIELDS[name]})


DexParams = _make_struct("DexParams")
DexSimViews = _make_struct("DexSimViews")
DexStateBuffers = _make_struct("DexStateBuffers")
DexOutBuffers = _make_struct("DexOutBuffers")
DexEpilogueBuffers = _make_struct("DexEpilogueBuffers")

OBS_KEY_NAMES = [None] * CONSTS["DEX_NUM_OBS_KEYS"]
for _k, _v in CONSTS.items():
    if _k.startswith("DEX_K_"):
        OBS_KEY_NAMES[_v] = _k[len("DEX_K_"):].lower()
REWARD_TERM_NAMES = [None] * CONSTS["DEX_NUM_REWARD_TERMS"]
for _k, _v in CONSTS.items():
    if _k.startswith("DEX_R_"):
        REWARD_TERM_NAMES[_v] = _k[len("DEX_R_"):].lower()
CRITERION_NAMES = [None] * CONSTS["DEX_NUM_CRITERIA"]
for _k, _v in CONSTS.items():
    if _k.startswith("DEX_C_"):
        CRITERION_NAMES[_v] = _k[len("DEX_C_"):].lower()


class DexStepError(RuntimeError):
    pass


_lib = None


def load(path: str = None) -> ctypes.CDLL:
    """Load libdexstep.so.  Raises (no fallback) if it is missing: build it with
    ``python -c 'import __graft_entry__ as g; g.build()'`` or ``python -m dexrobot_isaac_b200.build``."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise DexStepError(f"CUDA library not built: {path} is missing (run `python -m dexrobot_isaac_b200.build`). "
                           "There is no CPU or eager fallback for this path.")
    lib = ctypes.CDLL(path)
    lib.dex_strerror.restype = ctypes.c_char_p
    lib.dex_strerror.argtypes = [ctypes.c_int]
    for fn in FUNCTIONS:
        if not hasattr(lib, fn):
            raise DexStepError(f"{path} does not export {fn} (stale build?)")
    P = ctypes.POINTER
    i32, u32, u64, vp = ctypes.c_int32, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_void_p
    lib.dex_pre_physics.argtypes = [P(DexParams), P(DexStateBuffers), vp, vp, i32, vp]
    lib.dex_post_physics.argtypes = [P(DexParams), P(DexSimViews), P(DexStateBuffers), P(DexOutBuffers),
                                     P(DexEpilogueBuffers), i32, i32, u32, vp, u64, u64, vp]
    lib.dex_post_physics_noalloc.argtypes = lib.dex_post_physics.argtypes
    lib.dex_reset_actor_indices.argtypes = [P(DexEpilogueBuffers), i32, i32, i32, vp, vp, vp, vp]
    lib.dex_policy_input_update.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, vp, ctypes.c_float, ctypes.c_float, vp, vp]
    lib.dex_reset_idx.argtypes = [P(DexParams), P(DexSimViews), P(DexStateBuffers), vp, i32, vp, u64, u64, vp]
    lib.dex_step_epilogue.argtypes = [P(DexParams), P(DexEpilogueBuffers), i32, vp]
    lib.dex_apply_coupling.argtypes = [P(DexParams), vp, vp, i32, vp]
    lib.dex_extract_active_targets.argtypes = [P(DexParams), vp, vp, i32, vp]
    lib.dex_point_in_hand_frame.argtypes = [vp, vp, vp, vp, i32, vp]
    lib.dex_quat_mul.argtypes = [vp, vp, vp, i32, vp]
    lib.dex_debug_v2_launches.argtypes = []
    lib.dex_debug_launch_floor.argtypes = [i32, vp]
    lib.dex_debug_last_post_kernel.argtypes = []
    for fn in FUNCTIONS:
        if fn != "dex_strerror":
            getattr(lib, fn).restype = ctypes.c_int
    if lib.dex_abi_version() != CONSTS["DEX_ABI_VERSION"]:
        raise DexStepError(f"ABI mismatch: header {CONSTS['DEX_ABI_VERSION']} vs library {lib.dex_abi_version()}")
    if lib.dex_sizeof_params() != ctypes.sizeof(DexParams):
        raise DexStepError(f"DexParams layout mismatch: C {lib.dex_sizeof_params()} B vs ctypes {ctypes.sizeof(DexParams)} B")
    if path == LIB_PATH:
        _lib = lib
    return lib


def check(code: int, what: str = "dexstep call"):
    """Raise RuntimeError (the reference's fail-fast style) on a non-zero return code."""
    if code != 0:
        msg = load().dex_strerror(code).decode()
        raise DexStepError(f"{what} failed: {msg} (code {code})")
