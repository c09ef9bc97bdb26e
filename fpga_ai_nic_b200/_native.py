"""ctypes bindings to the in-tree native libraries.

``oracle()`` loads ``_lib/libainic_oracle.so`` (CPU; built on demand with g++).
``cuda()`` loads ``_lib/libainic_cuda.so`` (sm_100a kernels).  On a machine with a GPU a missing or
unloadable CUDA library is a hard error - ops never fall back to eager PyTorch silently.
"""
from __future__ import annotations

import ctypes as C
import os
from functools import lru_cache
from pathlib import Path

LIBDIR = Path(__file__).resolve().parent / "_lib"

c_f32p = C.POINTER(C.c_float)
c_u8p = C.POINTER(C.c_uint8)
c_i32p = C.POINTER(C.c_int32)
c_u64p = C.POINTER(C.c_uint64)
vp = C.c_void_p
sz = C.c_size_t


class NativeError(RuntimeError):
    pass


ERRORS = {0: "ok", -1: "bad argument", -2: "CUDA error", -3: "timeout", -4: "busy", -5: "aborted", -6: "unsupported"}


class GemmScatter(C.Structure):
    _fields_ = [("inbox", vp * 8), ("nranks", C.c_int), ("rank", C.c_int), ("shard_frames", C.c_uint),
                ("frame0", C.c_uint)]


class GemmEpilogue(C.Structure):
    _fields_ = [("bias", vp), ("mask_out", vp), ("mask_in", vp), ("alpha", C.c_float), ("relu", C.c_int),
                ("accumulate", C.c_int), ("cgrad", vp), ("codec", C.c_int), ("bfp_mode", C.c_int),
                ("scatter", C.POINTER(GemmScatter)), ("done_flag", vp), ("done_ticket", vp)]


class FlowChunk(C.Structure):
    _fields_ = [("offset", sz), ("count", sz), ("ready_index", C.c_int), ("ready_need", C.c_int)]


class AllreduceArgs(C.Structure):
    _fields_ = [("grad", vp), ("weight", vp), ("weight_bf16", vp), ("grad_out", vp), ("count", sz), ("lr", C.c_float),
                ("codec", C.c_int), ("bfp_mode", C.c_int), ("algo", C.c_int), ("max_ctas", C.c_int),
                ("grad_compressed", vp), ("grad_scattered", C.c_int)]


class Rqst(C.Structure):
    _fields_ = [("idx", C.c_int), ("count", C.c_int), ("buf", vp), ("dev_buf", vp), ("seq", C.c_uint64)]


def _ensure_built(name: str) -> Path:
    path = LIBDIR / name
    # build_*() are no-ops unless a source or header is newer than the library, so a stale .so left in
    # the tree after an edit to csrc/ is rebuilt instead of being loaded silently.  AINIC_NO_BUILD=1
    # (deployments without a compiler) trusts whatever is there.
    if path.exists() and os.environ.get("AINIC_NO_BUILD") == "1":
        return path
    from .utils import build
    force = bool(os.environ.get("AINIC_REBUILD"))
    try:
        if "oracle" in name:
            build.build_oracle(force=force)
        else:
            build.build_cuda(force=force)
    except Exception:
        if not path.exists():
            raise
        import warnings
        warnings.warn(f"could not rebuild {name}; loading the existing (possibly stale) library")
    return path


@lru_cache(maxsize=None)
def oracle() -> C.CDLL:
    lib = C.CDLL(str(_ensure_built("libainic_oracle.so")))
    lib.ainic_ref_bfp_bytes.restype = sz
    lib.ainic_ref_bfp_bytes.argtypes = [sz]
    lib.ainic_ref_mx_bytes.restype = sz
    lib.ainic_ref_mx_bytes.argtypes = [sz]
    lib.ainic_ref_bfp_compress.argtypes = [vp, sz, vp, C.c_int]
    lib.ainic_ref_bfp_decompress.argtypes = [vp, sz, vp, C.c_int]
    lib.ainic_ref_mx_compress.argtypes = [vp, sz, vp]
    lib.ainic_ref_mx_decompress.argtypes = [vp, sz, vp]
    lib.ainic_ref_bfp_roundtrip.argtypes = [vp, sz, C.c_int]
    lib.ainic_ref_allreduce_exact.argtypes = [vp, C.c_int, sz, vp]
    lib.ainic_ref_allreduce_ring.argtypes = [vp, C.c_int, sz, C.c_int, C.c_int, vp]
    lib.ainic_ref_allreduce_oneshot.argtypes = [vp, C.c_int, sz, C.c_int, C.c_int, vp]
    lib.ainic_ref_allreduce_twoshot.argtypes = [vp, C.c_int, sz, C.c_int, C.c_int, vp]
    lib.ainic_ref_sgd.argtypes = [vp, vp, sz, C.c_float]
    lib.ainic_ref_bucket_plan.restype = None
    lib.ainic_ref_bucket_plan.argtypes = [C.c_int, vp, sz, vp, C.c_int, vp, vp]
    lib.ainic_ref_fc_fwd.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.ainic_ref_fc_bwd.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.ainic_ref_softmax_loss.restype = C.c_float
    lib.ainic_ref_softmax_loss.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_float]
    lib.ainic_ref_init_buf.argtypes = [vp, sz, C.c_uint64, C.c_int, C.c_int]
    lib.ainic_ref_heap_new.restype = vp
    lib.ainic_ref_heap_new.argtypes = [sz]
    lib.ainic_ref_heap_delete.argtypes = [vp]
    lib.ainic_ref_heap_alloc.restype = sz
    lib.ainic_ref_heap_alloc.argtypes = [vp, sz, sz]
    lib.ainic_ref_heap_free.argtypes = [vp, sz]
    lib.ainic_ref_heap_stats.argtypes = [vp, vp]
    for fn, args, res in _NIC_SIGS:
        if hasattr(lib, fn):
            getattr(lib, fn).argtypes = args
            getattr(lib, fn).restype = res
    return lib


_NIC_SIGS = [
    ("ainic_nic_open", [C.POINTER(vp), vp, C.c_int, C.c_int, C.c_int], C.c_int),
    ("ainic_nic_close", [vp], C.c_int),
    ("ainic_nic_setup", [vp, sz, C.c_int, C.c_int, C.c_int, C.c_float], C.c_int),
    ("ainic_nic_done_buffer", [vp], C.POINTER(C.c_uint32)),
    ("ainic_nic_all_reduce", [vp, vp, vp, C.c_int, sz, C.POINTER(Rqst), vp], C.c_int),
    ("ainic_nic_wait", [vp, C.POINTER(Rqst), C.c_double], C.c_int),
    ("ainic_nic_abort", [vp], C.c_int),
    ("ainic_nic_latency_ns", [vp], C.c_uint64),
    ("ainic_nic_stall_ns", [vp], C.c_uint64),
    ("ainic_nic_fake_link", [C.POINTER(vp), C.c_int], C.c_int),
    ("ainic_iallreduce", [vp, vp, sz, C.POINTER(Rqst), vp], C.c_int),
]


def cuda_available() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@lru_cache(maxsize=None)
def cuda() -> C.CDLL:
    path = _ensure_built("libainic_cuda.so")
    try:
        lib = C.CDLL(str(path))
    except OSError as e:  # loud: never fall back silently on a GPU box
        raise NativeError(f"cannot load {path}: {e}") from e
    lib.ainic_version.restype = C.c_char_p
    lib.ainic_last_error.restype = C.c_char_p
    lib.ainic_bfp_bytes.restype = sz
    lib.ainic_bfp_bytes.argtypes = [sz]
    lib.ainic_mx_bytes.restype = sz
    lib.ainic_mx_bytes.argtypes = [sz]
    lib.ainic_bfp_compress.argtypes = [vp, sz, vp, C.c_int, vp]
    lib.ainic_bfp_decompress.argtypes = [vp, sz, vp, C.c_int, vp]
    lib.ainic_mx_compress.argtypes = [vp, sz, vp, vp]
    lib.ainic_mx_decompress.argtypes = [vp, sz, vp, vp]
    lib.ainic_sgd.argtypes = [vp, vp, vp, sz, C.c_float, vp]
    lib.ainic_cast_f32_bf16.argtypes = [vp, vp, sz, vp]
    lib.ainic_gemm_bf16.argtypes = [C.c_int, vp, C.c_int, vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.POINTER(GemmEpilogue), C.c_int, C.c_int, vp]
    lib.ainic_gemm_trace.argtypes = [vp, C.c_int]
    lib.ainic_gemm_tf32.argtypes = [C.c_int, vp, C.c_int, vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.POINTER(GemmEpilogue), C.c_int, C.c_int, vp]
    lib.ainic_mx_sf_bytes.restype = sz
    lib.ainic_mx_sf_bytes.argtypes = [sz, sz]
    lib.ainic_mx_quantize_rows.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp]
    lib.ainic_mx_quantize_cols.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp]
    lib.ainic_gemm_mxfp8.argtypes = [C.c_int, vp, C.c_int, vp, vp, C.c_int, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.POINTER(GemmEpilogue), C.c_int, C.c_int, vp]
    lib.ainic_softmax_xent.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_float, vp]
    lib.ainic_bias_grad.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp]
    lib.ainic_comm_create.argtypes = [C.POINTER(vp), C.c_int, C.c_int, sz]
    lib.ainic_comm_export.argtypes = [vp, vp]
    lib.ainic_comm_connect.argtypes = [vp, vp]
    lib.ainic_comm_destroy.argtypes = [vp]
    lib.ainic_comm_alloc.restype = vp
    lib.ainic_comm_alloc.argtypes = [vp, sz, sz]
    lib.ainic_comm_free.argtypes = [vp, vp]
    lib.ainic_comm_release_retired.argtypes = [vp]
    lib.ainic_comm_heap_stats.argtypes = [vp, vp]
    lib.ainic_comm_peer_ptr.restype = vp
    lib.ainic_comm_peer_ptr.argtypes = [vp, vp, C.c_int]
    lib.ainic_comm_counters.argtypes = [vp, c_u64p]
    lib.ainic_allreduce_fused.argtypes = [vp, C.POINTER(AllreduceArgs), vp]
    lib.ainic_allreduce_emulate.argtypes = [C.c_int, vp, vp, vp, vp, sz, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    lib.ainic_comm_flow_trace.argtypes = [vp, c_u64p]
    lib.ainic_allreduce_flow.argtypes = [vp, C.POINTER(AllreduceArgs), C.POINTER(FlowChunk), C.c_int, vp, vp]
    lib.ainic_allreduce_flow_emulate.argtypes = [C.c_int, vp, vp, vp, vp, sz, C.c_float, C.c_int, C.c_int, C.POINTER(sz),
                                                 C.c_int, C.c_int, vp]
    lib.ainic_nvls_supported.restype = C.c_int
    lib.ainic_nvls_stage1.argtypes = [C.POINTER(vp), C.c_int, C.c_int, sz, C.c_char_p]
    lib.ainic_nvls_stage2.argtypes = [vp]
    lib.ainic_nvls_stage3.argtypes = [vp]
    lib.ainic_nvls_alloc.restype = vp
    lib.ainic_nvls_alloc.argtypes = [vp, sz]
    lib.ainic_nvls_mc_ptr.restype = vp
    lib.ainic_nvls_mc_ptr.argtypes = [vp, vp]
    lib.ainic_nvls_destroy.argtypes = [vp]
    lib.ainic_allreduce_nvls.argtypes = [vp, vp, C.POINTER(AllreduceArgs), vp]
    lib.ainic_allreduce_emulate_noalloc.argtypes = [C.c_int, vp, vp, vp, vp, sz, C.c_float, C.c_int, C.c_int, C.c_int,
                                                    C.c_int, vp, C.c_char_p, sz]
    lib.ainic_mlp_create.argtypes = [C.POINTER(vp), vp, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int, C.c_float,
                                     C.c_float, C.c_int, C.c_int, C.c_int]
    lib.ainic_mlp_create2.argtypes = [C.POINTER(vp), vp, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_int, C.c_float,
                                      C.c_float, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.ainic_softmax_xent_f32.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_float, vp]
    lib.ainic_mlp_destroy.argtypes = [vp]
    lib.ainic_mlp_destroy_ex.argtypes = [vp, C.c_int]
    lib.ainic_mlp_buffer.restype = vp
    lib.ainic_mlp_buffer.argtypes = [vp, C.c_int, C.c_int]
    lib.ainic_mlp_sync_shadows.argtypes = [vp, vp]
    lib.ainic_mlp_forward.argtypes = [vp, vp, vp]
    lib.ainic_mlp_step.argtypes = [vp, vp, vp, vp, vp]
    lib.ainic_mlp_wait_updates.argtypes = [vp, vp]
    lib.ainic_mlp_profile.restype = C.c_ulonglong
    lib.ainic_mlp_profile.argtypes = [vp, C.c_int, C.POINTER(C.c_double)]
    lib.ainic_mlp_launches.restype = C.c_ulonglong
    lib.ainic_mlp_launches.argtypes = [vp]
    lib.ainic_mlp_set_bucket_bytes.restype = C.c_int
    lib.ainic_mlp_set_bucket_bytes.argtypes = [vp, C.c_size_t]
    lib.ainic_mlp_bucket_bytes.restype = C.c_size_t
    lib.ainic_mlp_bucket_bytes.argtypes = [vp]
    lib.ainic_mlp_loss_ptr.restype = vp
    lib.ainic_mlp_loss_ptr.argtypes = [vp]
    for fn, args, res in _NIC_SIGS:
        if hasattr(lib, fn):
            getattr(lib, fn).argtypes = args
            getattr(lib, fn).restype = res
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = ""
        try:
            msg = cuda().ainic_last_error().decode()
        except Exception:
            pass
        raise NativeError(f"{what or 'native call'} failed: {ERRORS.get(rc, rc)} {msg}".strip())


def stream_ptr(stream=None) -> int:
    """cudaStream_t of a torch stream (default: the current stream)."""
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream
