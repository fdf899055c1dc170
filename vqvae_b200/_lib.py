"""ctypes binding of libvq_b200.so (include/vq_b200.h).  No torch types cross this boundary:
only raw device pointers, sizes and a cudaStream_t handle.

The library is built in-tree by ``build.py``; if it is missing the import of the ops fails loudly --
there is no CPU or eager-PyTorch fallback behind it.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VQ_B200_LIB") or os.path.join(HERE, "libvq_b200.so")   # env: tuning variants only

VQ_OK, VQ_ERR_ARG, VQ_ERR_CUDA, VQ_ERR_UNSUPPORTED = 0, 1, 2, 3
VQ_DT_BF16, VQ_DT_F32, VQ_DT_F16 = 0, 1, 2
VQ_SEARCH_BF16, VQ_SEARCH_F32, VQ_SEARCH_F16 = 0, 1, 2
VQ_SEARCH_SIMT = 1024       # with VQ_SEARCH_F32: force the CUDA-core kernel
VQ_STAT_SUM_Z, VQ_STAT_SUM_RESID = 0, 1
VQ_STAT_ACCUMULATE = 16
VQ_STAT_PRESORTED = 32      # idx comes from vq_assign(... | VQ_SEARCH_PRESORT) on the same workspace
VQ_SEARCH_WS_CLEAN = 512    # caller vouches that the workspace counters are zero (see include/vq_b200.h)
VQ_SEARCH_PRESORT = 256     # vq_assign also leaves histogram / ranks / scan of idx in the workspace
ABI_VERSION = 5
MAX_PEERS, IPC_HANDLE_BYTES = 16, 64
NOALLOC_SCRATCH_BYTES = 4 << 20

_p, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float

# name -> (restype, argtypes); mirrors include/vq_b200.h one to one
SIGNATURES = {
    "vq_abi_version": (_i, []),
    "vq_last_error": (C.c_char_p, []),
    "vq_launch_count": (_i64, []),
    "vq_set_pdl_mask": (_i, [_i]),
    "vq_set_peer_timeout_ms": (_i64, [_i64]),
    "vq_norm_elems": (_i, [_i]),
    "vq_workspace_bytes": (_i, [_i64, _i, _i, _i, _i, C.POINTER(C.c_size_t)]),
    "vq_shadow_elems": (_i64, [_i, _i, _i]),
    "vq_prepare_codebook": (_i, [_p, _i, _i, _i, _p, _p, _p]),
    "vq_assign": (_i, [_p, _i, _i64, _i, _i64, _p, _p, _p, _i, _i, _p, _p, _p, _p]),
    "vq_gather_st_loss_scatter": (_i, [_p, _i, _i64, _i, _i64, _p, _p, _i, _p, _i, _i64, _p, _p, _p, _i, _i, _p, _p, _f,
                                       _p, _p]),
    "vq_gather_st_loss_scatter_noalloc": (_i, [_p, _i, _i64, _i, _i64, _p, _p, _i, _p, _i64, _p, _p, _p, _i, _p, _p, _f, _p]),
    "vq_ema_update": (_i, [_p, _p, _p, _p, _p, _i, _i, _f, _f, _i, _i, _p, _p, _p, _p, _p, _p, _i, _i, _p, _p]),
    "vq_backward": (_i, [_p, _p, _p, _i64, _i, _i64, _p, _f, _p, _i, _p]),
    "vq_codebook_grad": (_i, [_p, _i, _i, _p, _f, _p, _p]),
    "vq_cast_bf16": (_i, [_p, _i64, _i, _i64, _p, _p]),
    "vq_gather_rows_cf": (_i, [_p, _p, _i, _i, _i64, _i64, _p, _i, _p]),
    "vq_layout_transpose": (_i, [_p, _p, _i64, _i64, _i64, _i, _p]),
    "vq_ema_update_dp": (_i, [_p, _p, _p, _i, _i, _f, _f, _i, _i, _p, _p, _p, _p, _p, _p, _i, _i, _p,
                              C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), _i, _i, _i, _p, _p, _p, _p]),
    "vq_ipc_alloc": (_i, [C.c_size_t, C.POINTER(C.c_void_p), _p]),
    "vq_ipc_free": (_i, [_p]),
    "vq_ipc_open": (_i, [_p, C.POINTER(C.c_void_p)]),
    "vq_ipc_close": (_i, [_p]),
    "vq_knn_workspace_bytes": (_i, [_i64, _i, C.POINTER(C.c_size_t)]),
    "vq_knn_topk": (_i, [_p, _i64, _i, _p, _p, _f, _i, _p, _p, _p, _p]),
    "vq_knn_mutual_pairs": (_i, [_p, _i64, _i, _p, _p, _p, _p]),
}


class VQLibError(RuntimeError):
    pass


_lib = None


def load(build_if_missing: bool = False) -> C.CDLL:
    """dlopen the library (once) and set every prototype."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if build_if_missing:
            from . import build as _build
            _build.build()
        else:
            raise VQLibError(
                f"{LIB_PATH} is missing: build it with `python vqvae_b200/build.py` (nvcc, sm_100a). "
                "The vqvae_b200 ops have no fallback path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch
        fn.restype, fn.argtypes = res, args
    ver = lib.vq_abi_version()
    if ver != ABI_VERSION:
        raise VQLibError(f"libvq_b200.so ABI version {ver}, binding expects {ABI_VERSION}")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != VQ_OK:
        msg = load().vq_last_error().decode("utf-8", "replace")
        kind = {VQ_ERR_ARG: "bad argument", VQ_ERR_CUDA: "CUDA error", VQ_ERR_UNSUPPORTED: "unsupported"}.get(rc, rc)
        raise VQLibError(f"{what}: {kind}: {msg}")
