"""ctypes binding of ``liblmi_b200.so`` (the C ABI declared in ``include/lmi_b200.h``).

The library is built in-tree by ``build_library()`` (nvcc, sm_100a only).  There is
no CPU fallback: ``lib()`` raises if the shared object is missing or cannot be loaded.
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess
from ctypes import c_char_p, c_int, c_int64, c_void_p

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC_DIR = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "liblmi_b200.so")
SOURCES = ["lmi_mlp.cu", "lmi_mlp_tc.cu", "lmi_invert.cu", "lmi_scan.cu", "lmi_merge.cu", "lmi_layout.cu", "lmi_navigate.cu", "lmi_nav_fused.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
]

_lib = None


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.isfile(cand):
            return cand
    raise RuntimeError("nvcc not found; cannot build liblmi_b200.so")


def library_is_stale() -> bool:
    if not os.path.isfile(LIB_PATH):
        return True
    built = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC_DIR, f) for f in os.listdir(CSRC_DIR)]
    deps.append(os.path.join(os.path.dirname(PKG_DIR), "include", "lmi_b200.h"))
    return any(os.path.getmtime(p) > built for p in deps if os.path.isfile(p))


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a into ``liblmi_b200.so`` (in-tree)."""
    if not force and not library_is_stale():
        return LIB_PATH
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + SOURCES
    proc = subprocess.run(cmd, cwd=CSRC_DIR, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError(f"nvcc failed ({proc.returncode}):\n{proc.stdout}\n{proc.stderr}")
    if verbose:
        print(proc.stderr)
    return LIB_PATH


_SIGNATURES = {
    # name: (restype, argtypes)
    "lmi_version": (c_int, []),
    "lmi_error_string": (c_char_p, [c_int]),
    "lmi_linear": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "lmi_split_tf32": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
    "lmi_linear_tc": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                              c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "lmi_linear_tc_grouped": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_int, c_int, c_void_p]),
    "lmi_softmax_segments": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_void_p, c_int, c_void_p]),
    "lmi_softmax_topn": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "lmi_mlp_topn_supported": (c_int, [c_int, c_int, c_int, c_int, c_int]),
    "lmi_mlp_topn_workspace_bytes": (c_int64, [c_int, c_int, c_int]),
    "lmi_mlp_topn": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p,
                             c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                             c_void_p]),
    "lmi_argmax": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p]),
    "lmi_softmax_probs": (c_int, [c_void_p, c_int, c_int, c_void_p, c_int64, c_int64, c_void_p]),
    "lmi_navigate_tree": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p,
                                  c_void_p]),
    "lmi_scatter_rows": (c_int, [c_void_p, c_int, c_void_p, c_int64, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "lmi_cast_f16": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "lmi_max_items": (c_int64, [c_int, c_int, c_int, c_int]),
    "lmi_invert_workspace_ints": (c_int64, [c_int, c_int, c_int]),
    "lmi_workspace_layout": (c_int64, [c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "lmi_workspace_bytes": (c_int64, [c_int, c_int, c_int, c_int]),
    "lmi_invert": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                           c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "lmi_invert_static": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]),
    "lmi_scan_topk": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p,
                              c_int64, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "lmi_merge_topk": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                               c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "lmi_merge_ranks": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_int,
                                c_void_p, c_void_p, c_void_p]),
    "lmi_gather_peer": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p]),
    "lmi_merge_ranks_peer": (c_int, [c_void_p, c_int64, c_int64, c_int, c_int, c_int, c_void_p, c_void_p,
                                     c_void_p]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def lib() -> ctypes.CDLL:
    """The loaded library.  Raises (never falls back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(learnedmetricindex_b200 has no CPU fallback)")
    handle = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(handle, name)
        fn.restype = res
        fn.argtypes = args
    _lib = handle
    return _lib


class LmiError(RuntimeError):
    pass


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = lib().lmi_error_string(code)
        raise LmiError(f"{what or 'lmi call'} failed with code {code}: {msg.decode() if msg else '?'}")
