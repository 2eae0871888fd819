"""ctypes binding of libhg_b200.so (the C ABI declared in include/hg_b200.h).

PyTorch is used for device memory and streams only; every compute call goes
through the C ABI with raw device pointers.  There is NO CPU fallback: if the
shared library is missing or no CUDA device is present the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# HG_B200_LIB: load another build of the same library (A/B runs of two kernel variants on one box)
LIB_PATH = os.environ.get("HG_B200_LIB") or os.path.join(_HERE, "libhg_b200.so")

HG_OK, HG_ERR_BAD_ARG, HG_ERR_UNSUPPORTED, HG_ERR_CUDA, HG_ERR_WORKSPACE = 0, -1, -2, -3, -4
METRIC = {"euclidean": 0, "gaussian": 1, "cosine": 2}
FORM_EXPANSION, FORM_DIRECT = 0, 1

_i64, _i32, _f32, _vp, _sz = C.c_int64, C.c_int, C.c_float, C.c_void_p, C.c_size_t
_psz = C.POINTER(C.c_size_t)
_pi64 = C.POINTER(C.c_int64)

class KnnPart(C.Structure):
    """struct hg_knn_part of include/hg_b200.h"""
    _fields_ = [("n", _i64), ("d", _i64), ("row_begin", _i64), ("row_end", _i64), ("reserved0", _i64),
                ("pos", _vp), ("pos_dim", C.c_int32), ("metric", C.c_int32), ("k", C.c_int32),
                ("exclude_self", C.c_int32), ("reserved1", C.c_int32), ("reserved2", C.c_int32),
                ("lambda_h", _f32), ("lambda_g", _f32)]


_pst = C.POINTER(KnnPart)

# name -> (restype, argtypes); kept in one table so tests can check that every
# symbol of include/hg_b200.h is exported
SIGNATURES = {
    "hg_version": (_i32, []),
    "hg_last_error": (C.c_char_p, []),
    "hg_launch_count": (_i64, []),
    "hg_profile_enable": (_i32, [_i32]),
    "hg_profile_collect": (_i32, [C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "hg_operand_bytes": (_sz, [_i64, _i64]),
    "hg_split_operands": (_i32, [_vp, _i64, _i64, _vp, _vp, _vp]),
    "hg_knn_feature_workspace_bytes": (_i32, [_i64, _i64, _i32, _i64, _i64, _psz]),
    "hg_knn_feature_plan": (_i32, [_i64, _i64, _i32, _i64, _i64, C.POINTER(C.c_int), C.POINTER(C.c_double)]),
    "hg_knn_feature_products": (_i32, [_i64, _i64, _i32, _i64, _i64, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "hg_knn_feature": (_i32, [_vp, _i64, _i64, _vp, _i32, _i32, _f32, _f32, _i32, _i64, _i64, _i32, _i32,
                              _vp, _vp, _vp, _sz, _vp]),
    "hg_knn_feature_cos": (_i32, [_vp, _i64, _i64, _vp, _i32, _i32, _f32, _f32, _i32, _i64, _i64, _i32, _i32,
                                  _vp, _vp, _vp, _vp, _sz, _vp]),
    "hg_edge_weights_from_knn": (_i32, [_vp, _i64, _i64, _vp, _vp, _i32, _vp, _i64, _vp, _vp]),
    "hg_similarity_dense": (_i32, [_vp, _i64, _vp, _i64, _i64, _vp, _vp, _i32, _f32, _f32, _i32, _vp, _i64, _vp]),
    "hg_stats_workspace_bytes": (_i32, [_i64, _psz]),
    "hg_similarity_stats": (_i32, [_vp, _i64, _i64, _vp, _vp, _sz, _vp]),
    "hg_similarity_dense_tc_workspace_bytes": (_i32, [_i64, _i64, _psz]),
    "hg_similarity_dense_tc": (_i32, [_vp, _i64, _i64, _vp, _i32, _f32, _f32, _vp, _i64, _vp, _sz, _vp]),
    "hg_similarity_stats_streaming_workspace_bytes": (_i32, [_i64, _i64, _psz]),
    "hg_similarity_stats_streaming": (_i32, [_vp, _i64, _i64, _vp, _i32, _f32, _f32, _vp, _vp, _sz, _vp]),
    "hg_knn_spatial_workspace_bytes": (_i32, [_i64, _i32, _i64, _i64, _psz]),
    "hg_knn_spatial": (_i32, [_vp, _i64, _i32, _i32, _i64, _i64, _vp, _vp, _vp, _sz, _vp]),
    "hg_edges_workspace_bytes": (_i32, [_i64, _i64, _i32, _i64, _psz]),
    "hg_clique_pairs": (_i32, [_vp, _i64, _i32, _pi64, _vp]),
    "hg_edges_count": (_i32, [C.POINTER(_vp), C.POINTER(_i32), _i32, _i64, _vp, _i32, _i64, _pi64, _vp, _sz, _vp]),
    "hg_edges_fill": (_i32, [_i64, _vp, _vp, _sz, _vp]),
    "hg_edges_build": (_i32, [_i64, _vp, _vp, _sz, _vp]),
    "hg_edge_weights_cosine": (_i32, [_vp, _i64, _i64, _vp, _i64, _vp, _vp]),
    "hg_csr_workspace_bytes": (_i32, [_i64, _i64, _psz]),
    "hg_csr_from_coo": (_i32, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _sz, _vp]),
    "hg_csr_from_coo_static": (_i32, [_vp, _i64, _i64, _vp, _vp, _vp, _vp]),
    "hg_threshold_workspace_bytes": (_i32, [_i64, _psz]),
    "hg_threshold_edges_count": (_i32, [_vp, _i64, _f32, _pi64, _vp, _sz, _vp]),
    "hg_kmeans_assign": (_i32, [_vp, _i64, _i64, _vp, _i32, _vp, _vp, _vp]),
    "hg_kmeans_workspace_bytes": (_i32, [_i64, _i32, _psz]),
    "hg_kmeans_update": (_i32, [_vp, _i64, _i64, _vp, _i32, _vp, _vp, _vp, _vp, _sz, _vp]),
    "hg_kmeans_inertia": (_i32, [_vp, _i64, _vp, _vp]),
    "hg_kmeans_candidate_d2": (_i32, [_vp, _i64, _i64, _vp, _i32, _vp, _vp]),
    "hg_kmeans_fit_workspace_bytes": (_i32, [_i64, _i64, _i32, _psz]),
    "hg_kmeans_fit": (_i32, [_vp, _i64, _i64, _i32, _i32, _i32, C.c_double, C.c_uint64, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "hg_build_workspace_bytes": (_i32, [_i64, _i64, _i32, _i32, _i32, _i64, _i64, _i32, _psz]),
    "hg_build_knn_hypergraph": (_i32, [_vp, _i64, _i64, _vp, _i32, _i32, _f32, _f32, _i32, _i32, _vp, _i32, _i64,
                                       _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _pi64, _vp, _sz, _vp]),
    "hg_build_knn_hypergraph_enqueue": (_i32, [_vp, _i64, _i64, _vp, _i32, _i32, _f32, _f32, _i32, _i32, _vp, _i32, _i64,
                                               _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "hg_build_batch_workspace_bytes": (_i32, [_vp, _i32, _i64, _i32, _i32, _i32, _psz]),
    "hg_build_batch_enqueue": (_i32, [_vp, _i32, _i64, _i32, _i32, _f32, _f32, _i32, _i32, _vp, _sz, _vp, _i32, _i32]),
    "hg_build_batch_grouped_workspace_bytes": (_i32, [_vp, _i32, _i64, _i32, _i32, _i32, _i64, _psz, _psz]),
    "hg_build_batch_grouped_enqueue": (_i32, [_vp, _i32, _i64, _i32, _i32, _f32, _f32, _i32, _i32, _i64,
                                              _vp, _sz, _vp, _vp, _sz, _vp, _i32]),
    "hg_hypergraph_propagate": (_i32, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp]),
    "hg_enable_peer_access": (_i32, [_i32]),
    "hg_ipc_open": (_i32, [C.c_char_p, C.POINTER(_vp)]),
    "hg_ipc_close": (_i32, [_vp]),
    "hg_peer_alloc": (_i32, [_sz, C.POINTER(_vp), C.c_char_p]),
    "hg_peer_free": (_i32, [_vp]),
    "hg_knn_sympart_workspace_bytes": (_i32, [_pst, _psz]),
    "hg_knn_sympart_prepass": (_i32, [_pst, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "hg_knn_sympart_sweep": (_i32, [_pst, _vp, _i64, _i32, _vp, _vp, _vp, _vp, _sz, _vp]),
    "hg_knn_sympart_finish": (_i32, [_pst, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "hg_aggregate_workspace_bytes": (_i32, [_i64, _i64, _i32, _psz]),
    "hg_aggregate_super_patches": (_i32, [_vp, _i64, _i64, _vp, _i32, _vp, _i32, _f32, _f32, _vp, _i64, _i32,
                                          _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "hg_threshold_edges_fill": (_i32, [_vp, _i64, _f32, _i64, _vp, _vp, _vp, _sz, _vp]),
    "hg_threshold_edges_build": (_i32, [_vp, _i64, _f32, _i64, _vp, _vp, _vp, _sz, _vp]),
}

_lib = None


def load_library() -> C.CDLL:
    """dlopen the in-tree shared library (built by `__graft_entry__.build()` / make)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def last_error() -> str:
    return load_library().hg_last_error().decode("utf-8", "replace")


def check(rc: int) -> None:
    """Map C ABI return codes onto the exception types the reference raises."""
    if rc == HG_OK:
        return
    msg = last_error()
    if rc in (HG_ERR_BAD_ARG, HG_ERR_UNSUPPORTED):
        raise ValueError(msg)          # reference: ValueError for N<=1, bad method, ... (similarity_kernel.py:176)
    raise RuntimeError(f"hg_b200 error {rc}: {msg}")


def require_cuda(device=None) -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("hg_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError(f"hg_b200 runs on CUDA devices only, got device={dev}")
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    return dev


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


def stream_ptr(dev) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


def f32c(t: torch.Tensor, dev) -> torch.Tensor:
    """float32, contiguous, on `dev` (host tensors are copied through pinned staging by torch)."""
    return t.to(device=dev, dtype=torch.float32, non_blocking=True).contiguous()


class Workspace:
    """Grow-only device scratch buffer, one per device; 1024-byte aligned."""
    _cache = {}

    @classmethod
    def get(cls, dev, nbytes: int, slot: str = "main") -> torch.Tensor:
        # one buffer per (device, stream, purpose): builds running on different streams never share scratch
        key = (dev.index, torch.cuda.current_stream(dev).cuda_stream, slot)
        buf = cls._cache.get(key)
        if buf is None or buf.numel() < nbytes:
            buf = None
            cls._cache.pop(key, None)
            buf = torch.empty(max(nbytes, 1 << 20) + 1024, dtype=torch.uint8, device=dev)
            cls._cache[key] = buf
        off = (-buf.data_ptr()) % 1024
        return buf[off:off + max(nbytes, 1)]

    @classmethod
    def clear(cls):
        cls._cache.clear()
