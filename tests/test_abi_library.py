"""CPU tests of the C-ABI boundary itself (no GPU, no kernel launch): libdexstep.so loads, exports every function that
include/dexstep.h declares, reports the header's ABI version and struct size, and its argument validation - which runs before
anything touches CUDA - answers with the documented error codes."""
import ctypes
import os
import re

import pytest

from dexrobot_isaac_b200 import _lib as L
from dexrobot_isaac_b200 import build as B

HEADER = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "dexstep.h")


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(B.LIB_PATH):
        B.build()  # nvcc cross-compiles for sm_100a without a GPU
    return L.load()


def declared_functions():
    """Every `int dex_*(` / `const char* dex_*(` prototype of the header, found independently of _lib's own parser."""
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    return sorted(set(re.findall(r"^(?:int|const char\*)\s+(dex_\w+)\s*\(", text, flags=re.M)))


def test_header_declares_the_path_and_lib_parser_agrees():
    names = declared_functions()
    for must in ("dex_pre_physics", "dex_post_physics", "dex_post_physics_noalloc", "dex_reset_idx", "dex_step_epilogue",
                 "dex_reset_actor_indices", "dex_policy_input_update", "dex_strerror", "dex_abi_version", "dex_sizeof_params"):
        assert must in names
    assert sorted(L.FUNCTIONS) == names


def test_library_exports_every_declared_symbol(lib):
    raw = ctypes.CDLL(B.LIB_PATH)  # a second, plain handle: no argtypes, nothing from _lib
    missing = [n for n in declared_functions() if not hasattr(raw, n)]
    assert not missing, f"libdexstep.so does not export {missing}"


def test_abi_version_and_struct_size(lib):
    assert lib.dex_abi_version() == L.CONSTS["DEX_ABI_VERSION"]
    assert lib.dex_sizeof_params() == ctypes.sizeof(L.DexParams)
    assert "rigid_row_floats" in [n for n, _ in L.DexParams._fields_]


def test_strerror_names_every_library_code(lib):
    codes = {k: v for k, v in L.CONSTS.items() if k.startswith("DEX_ERR_")}
    assert "DEX_ERR_CAPACITY" in codes and all(v < 0 for v in codes.values())
    texts = {k: lib.dex_strerror(v).decode() for k, v in codes.items()}
    assert all(t and "unknown" not in t.lower() for t in texts.values()), texts
    assert len(set(texts.values())) == len(texts)
    assert lib.dex_strerror(0).decode().lower() in ("ok", "success", "no error")


def test_validation_runs_before_any_launch(lib):
    """NULL / stale parameter blocks are refused with library codes; none of these calls reaches CUDA (this box has no GPU)."""
    C = L.CONSTS
    assert lib.dex_pre_physics(None, None, None, None, 64, None) == C["DEX_ERR_NULL"]
    assert lib.dex_post_physics(None, None, None, None, None, 64, 0, 0, None, 0, 0, None) == C["DEX_ERR_NULL"]
    P = L.DexParams()
    P.abi_version = C["DEX_ABI_VERSION"] - 1
    st = L.DexStateBuffers()
    assert lib.dex_pre_physics(ctypes.byref(P), ctypes.byref(st), None, None, 64, None) == C["DEX_ERR_VERSION"]
    P.abi_version = C["DEX_ABI_VERSION"]
    P.task = 99
    assert lib.dex_pre_physics(ctypes.byref(P), ctypes.byref(st), None, None, 64, None) == C["DEX_ERR_PARAM"]
