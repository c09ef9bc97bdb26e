"""Front-ends of the hand-written sm_100a kernels (torch CUDA tensors in/out, launched on the
current torch stream).  No op here has a PyTorch fallback: if ``libainic_cuda.so`` is missing the
import of the native library raises.

Reference counterparts: codec = hw/bfp_adapter.sv (+ cores), sgd = hw/weight_update.sv:440-452,
gemm_* = libxsmm_dnn_fc_{fwd,bwd}_exec_f32 (sw/mlp_mpi_example_f32.cpp:708-709, :741-742),
softmax_xent = libxsmm_dnn_smax_{fwd,bwd}_exec_f32 (:718-728).
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native as N

HW_EXACT, CLEAN = 0, 1
CODEC_F32, CODEC_BFP, CODEC_MXFP8, CODEC_BF16 = 0, 1, 2, 3
ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_RING, ALGO_NVLS = 0, 1, 2, 3, 4
GEMM_NT, GEMM_NN, GEMM_TN = 0, 1, 2

CODEC_NAMES = {"f32": CODEC_F32, "bfp": CODEC_BFP, "mxfp8": CODEC_MXFP8, "bf16": CODEC_BF16}
ALGO_NAMES = {"auto": ALGO_AUTO, "oneshot": ALGO_ONESHOT, "twoshot": ALGO_TWOSHOT, "ring": ALGO_RING, "nvls": ALGO_NVLS}

# launches of our own kernels since the last reset (bench.py reports it as gpu_launches)
launch_count = 0


def _count(n: int = 1) -> None:
    global launch_count
    launch_count += n


def _chk(t: torch.Tensor, dtype) -> torch.Tensor:
    assert t.is_cuda and t.dtype == dtype and t.is_contiguous(), (t.device, t.dtype, t.is_contiguous())
    return t


# ---------------------------------------------------------------- codecs
def bfp_bytes(count: int) -> int:
    return (count + 511) // 512 * 544


def mx_bytes(count: int) -> int:
    return (count + 511) // 512 * 528


def bfp_compress(x: torch.Tensor, mode: int = HW_EXACT, out: torch.Tensor | None = None) -> torch.Tensor:
    _chk(x, torch.float32)
    n = x.numel()
    if out is None:
        out = torch.empty(bfp_bytes(n), dtype=torch.uint8, device=x.device)
    N.check(N.cuda().ainic_bfp_compress(x.data_ptr(), n, out.data_ptr(), mode, N.stream_ptr()), "bfp_compress")
    _count()
    return out


def bfp_decompress(buf: torch.Tensor, count: int, mode: int = HW_EXACT, out: torch.Tensor | None = None) -> torch.Tensor:
    _chk(buf, torch.uint8)
    if out is None:
        out = torch.empty(count, dtype=torch.float32, device=buf.device)
    N.check(N.cuda().ainic_bfp_decompress(buf.data_ptr(), count, out.data_ptr(), mode, N.stream_ptr()), "bfp_decompress")
    _count()
    return out


def mx_compress(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    _chk(x, torch.float32)
    n = x.numel()
    if out is None:
        out = torch.empty(mx_bytes(n), dtype=torch.uint8, device=x.device)
    N.check(N.cuda().ainic_mx_compress(x.data_ptr(), n, out.data_ptr(), N.stream_ptr()), "mx_compress")
    _count()
    return out


def mx_decompress(buf: torch.Tensor, count: int, out: torch.Tensor | None = None) -> torch.Tensor:
    _chk(buf, torch.uint8)
    if out is None:
        out = torch.empty(count, dtype=torch.float32, device=buf.device)
    N.check(N.cuda().ainic_mx_decompress(buf.data_ptr(), count, out.data_ptr(), N.stream_ptr()), "mx_decompress")
    _count()
    return out


# ---------------------------------------------------------------- optimizer
def sgd_(w: torch.Tensor, g: torch.Tensor, lr: float, w_bf16: torch.Tensor | None = None) -> None:
    """In place: w = fmaf(-lr, g, w); optionally refresh the bf16 shadow."""
    _chk(w, torch.float32)
    _chk(g, torch.float32)
    assert w.numel() == g.numel()
    sh = _chk(w_bf16, torch.bfloat16).data_ptr() if w_bf16 is not None else None
    N.check(N.cuda().ainic_sgd(w.data_ptr(), g.data_ptr(), sh, w.numel(), lr, N.stream_ptr()), "sgd")
    _count()


def cast_bf16(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    _chk(x, torch.float32)
    if out is None:
        out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    N.check(N.cuda().ainic_cast_f32_bf16(x.data_ptr(), out.data_ptr(), x.numel(), N.stream_ptr()), "cast")
    _count()
    return out


# ---------------------------------------------------------------- GEMMs
def _gemm(layout, a, b, c, M, Nn, K, epi: N.GemmEpilogue | None, block_n=0, max_ctas=0, stream=None):
    lib = N.cuda()
    rc = lib.ainic_gemm_bf16(layout, a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0), c.data_ptr(), c.stride(0),
                             M, Nn, K, C.byref(epi) if epi is not None else None, block_n, max_ctas,
                             N.stream_ptr(stream))
    N.check(rc, "gemm_bf16")
    _count()


def gemm_fwd(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None = None, relu: bool = False,
             mask_out: torch.Tensor | None = None, out: torch.Tensor | None = None, block_n: int = 0,
             max_ctas: int = 0) -> torch.Tensor:
    """y[mb,n] = act(x[mb,k] @ w[n,k]^T + bias)  (bf16 in/out, fp32 accumulate)."""
    _chk(x, torch.bfloat16)
    _chk(w, torch.bfloat16)
    mb, k = x.shape
    n = w.shape[0]
    assert w.shape[1] == k
    if out is None:
        out = torch.empty(mb, n, dtype=torch.bfloat16, device=x.device)
    epi = N.GemmEpilogue(bias.data_ptr() if bias is not None else None,
                         mask_out.data_ptr() if mask_out is not None else None, None, 1.0, int(relu), 0)
    _gemm(GEMM_NT, x, w, out, mb, n, k, epi, block_n, max_ctas)
    return out


def gemm_bwd_data(dy: torch.Tensor, w: torch.Tensor, mask_in: torch.Tensor | None = None,
                  out: torch.Tensor | None = None, block_n: int = 0, max_ctas: int = 0) -> torch.Tensor:
    """dx[mb,k] = dy[mb,n] @ w[n,k], zeroed where mask_in == 0 (ReLU backward of the previous layer)."""
    _chk(dy, torch.bfloat16)
    _chk(w, torch.bfloat16)
    mb, n = dy.shape
    k = w.shape[1]
    assert w.shape[0] == n
    if out is None:
        out = torch.empty(mb, k, dtype=torch.bfloat16, device=dy.device)
    epi = N.GemmEpilogue(None, None, mask_in.data_ptr() if mask_in is not None else None, 1.0, 0, 0)
    _gemm(GEMM_NN, dy, w, out, mb, k, n, epi, block_n, max_ctas)
    return out


def gemm_bwd_weight(dy: torch.Tensor, x: torch.Tensor, out: torch.Tensor | None = None, accumulate: bool = False,
                    alpha: float = 1.0, block_n: int = 0, max_ctas: int = 0, stream=None,
                    compressed_out: torch.Tensor | None = None, codec: int = CODEC_BFP, bfp_mode: int = CLEAN,
                    scatter: tuple | None = None) -> torch.Tensor:
    """dw[n,k] = dy[mb,n]^T @ x[mb,k] as fp32, written straight into the flat gradient bucket.

    With ``compressed_out`` (uint8, ``bfp_bytes(n*k)`` / ``mx_bytes(n*k)``) the epilogue instead emits the
    gradient already in BFP / MXFP8 frame layout - what ``Communicator.allreduce_fused(grad_compressed=...)``
    consumes - and no fp32 gradient is written.

    ``scatter=(inboxes, rank, shard_frames, frame0)`` turns the epilogue into a reduce-scatter: frame f of
    the bucket (this GEMM's frames start at ``frame0``) is stored into ``inboxes[f // shard_frames]`` at slot
    ``[rank][f % shard_frames]`` (ainic_gemm_scatter_t).  ``inboxes`` are uint8 tensors or raw pointers."""
    _chk(dy, torch.bfloat16)
    _chk(x, torch.bfloat16)
    mb, n = dy.shape
    k = x.shape[1]
    assert x.shape[0] == mb
    if scatter is not None:
        inboxes, rank, shard_frames, frame0 = scatter
        sc = N.GemmScatter()
        for i, t in enumerate(inboxes):
            sc.inbox[i] = t.data_ptr() if isinstance(t, torch.Tensor) else int(t)
        sc.nranks, sc.rank, sc.shard_frames, sc.frame0 = len(inboxes), rank, shard_frames, frame0
        epi = N.GemmEpilogue(None, None, None, alpha, 0, 0, None, codec, bfp_mode, C.pointer(sc))
        rc = N.cuda().ainic_gemm_bf16(GEMM_TN, dy.data_ptr(), dy.stride(0), x.data_ptr(), x.stride(0), None, k, n, k, mb,
                                      C.byref(epi), block_n, max_ctas, N.stream_ptr(stream))
        N.check(rc, "gemm_bf16(scatter)")
        _count()
        return None
    if compressed_out is not None:
        _chk(compressed_out, torch.uint8)
        epi = N.GemmEpilogue(None, None, None, alpha, 0, 0, compressed_out.data_ptr(), codec, bfp_mode)
        lib = N.cuda()
        rc = lib.ainic_gemm_bf16(GEMM_TN, dy.data_ptr(), dy.stride(0), x.data_ptr(), x.stride(0), None, k, n, k, mb,
                                 C.byref(epi), block_n, max_ctas, N.stream_ptr(stream))
        N.check(rc, "gemm_bf16(compressed)")
        _count()
        return compressed_out
    if out is None:
        out = torch.empty(n, k, dtype=torch.float32, device=dy.device)
    epi = N.GemmEpilogue(None, None, None, alpha, 0, int(accumulate))
    _gemm(GEMM_TN, dy, x, out.view(n, k), n, k, mb, epi, block_n, max_ctas, stream)
    return out


# ---------------------------------------------------------------- loss
def softmax_xent(logits: torch.Tensor, labels: torch.Tensor, loss_weight: float = 0.1, want_probs: bool = False,
                 dlogits: torch.Tensor | None = None, loss_sum: torch.Tensor | None = None):
    """Returns (loss_sum[1] fp32 = sum of per-row CE, dlogits bf16, probs bf16 | None)."""
    _chk(logits, torch.bfloat16)
    _chk(labels, torch.int32)
    mb, n = logits.shape
    if dlogits is None:
        dlogits = torch.empty_like(logits)
    if loss_sum is None:
        loss_sum = torch.zeros(1, dtype=torch.float32, device=logits.device)
    probs = torch.empty_like(logits) if want_probs else None
    N.check(N.cuda().ainic_softmax_xent(logits.data_ptr(), labels.data_ptr(), probs.data_ptr() if want_probs else None,
                                        dlogits.data_ptr(), loss_sum.data_ptr(), mb, n, logits.stride(0), loss_weight,
                                        N.stream_ptr()), "softmax_xent")
    _count()
    return loss_sum, dlogits, probs


def bias_grad(dy: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    _chk(dy, torch.bfloat16)
    mb, n = dy.shape
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=dy.device)
    N.check(N.cuda().ainic_bias_grad(dy.data_ptr(), out.data_ptr(), mb, n, dy.stride(0), N.stream_ptr()), "bias_grad")
    _count(2)
    return out


# ---------------------------------------------------------------- multi-rank emulation on one GPU
def allreduce_emulate(grads: torch.Tensor, weights: torch.Tensor | None, lr: float, codec: int, algo: int,
                      bfp_mode: int = HW_EXACT, want_shadow: bool = False, want_grads: bool = True,
                      ctas_per_rank: int = 4):
    """Run an N-rank fused all-reduce(+SGD) with all ranks' buffers on one GPU (one cooperative launch).

    grads: [nranks, count] fp32 (fp32 in-place algorithms overwrite it), weights: [nranks, count] or None.
    Returns (grads_out [nranks, count] | None, weights (updated in place), shadow bf16 | None).
    """
    _chk(grads, torch.float32)
    nranks, count = grads.shape
    gout = torch.zeros_like(grads) if want_grads else None
    shadow = torch.zeros(nranks, count, dtype=torch.bfloat16, device=grads.device) if want_shadow else None
    if weights is not None:
        _chk(weights, torch.float32)
    rc = N.cuda().ainic_allreduce_emulate(nranks, grads.data_ptr(), weights.data_ptr() if weights is not None else None,
                                          shadow.data_ptr() if shadow is not None else None,
                                          gout.data_ptr() if gout is not None else None, count, lr, codec, bfp_mode,
                                          algo, ctas_per_rank, N.stream_ptr())
    N.check(rc, "allreduce_emulate")
    _count()
    return gout, weights, shadow


def allreduce_flow_emulate(grads: torch.Tensor, weights: torch.Tensor | None, lr: float, codec: int, chunk_counts,
                           bfp_mode: int = HW_EXACT, want_shadow: bool = False, want_grads: bool = True,
                           ctas_per_rank: int = 4):
    """The chunk-pipelined (flow) all-reduce with N ranks emulated on one GPU.  ``chunk_counts`` cut the
    bucket into consecutive chunks (multiples of 512 values) that are processed last chunk first, the
    way a backward pass produces its layers.  Same return values as :func:`allreduce_emulate`."""
    _chk(grads, torch.float32)
    nranks, count = grads.shape
    gout = torch.zeros_like(grads) if want_grads else None
    shadow = torch.zeros(nranks, count, dtype=torch.bfloat16, device=grads.device) if want_shadow else None
    if weights is not None:
        _chk(weights, torch.float32)
    cc = (N.sz * len(chunk_counts))(*[int(c) for c in chunk_counts])
    rc = N.cuda().ainic_allreduce_flow_emulate(nranks, grads.data_ptr(), weights.data_ptr() if weights is not None else None,
                                               shadow.data_ptr() if shadow is not None else None,
                                               gout.data_ptr() if gout is not None else None, count, lr, codec, bfp_mode,
                                               cc, len(chunk_counts), ctas_per_rank, N.stream_ptr())
    N.check(rc, "allreduce_flow_emulate")
    _count()
    return gout, weights, shadow


def gemm_trace(buf: torch.Tensor | None, launches: int = 0) -> None:
    """Diagnostics (ainic_gemm_trace): the next ``launches`` bf16 GEMM launches write per-CTA time stamps into
    ``buf`` (int64, ``[launches, 148, 8]``, zero-filled by the caller); ``None`` switches it off."""
    if buf is not None:
        assert buf.dtype == torch.int64 and buf.is_cuda and buf.is_contiguous() and buf.numel() >= launches * 148 * 8
    N.check(N.cuda().ainic_gemm_trace(buf.data_ptr() if buf is not None else None, int(launches)), "gemm_trace")


# ---------------------------------------------------------------- tf32 and MX block-scaled GEMMs (csrc/gemm/gemm_kinds.cu)
def gemm_tf32(layout: int, a: torch.Tensor, b: torch.Tensor, out: torch.Tensor | None = None, alpha: float = 1.0,
              accumulate: bool = False, block_n: int = 0, max_ctas: int = 0) -> torch.Tensor:
    """fp32 operands on the tensor cores as tf32, fp32 result.  Layouts as the bf16 GEMM:
    GEMM_NT  c[M,N] = a[M,K] @ b[N,K]^T;  GEMM_NN  c[M,N] = a[M,K] @ b[K,N];  GEMM_TN  c[M,N] = a[K,M]^T @ b[K,N]."""
    _chk(a, torch.float32)
    _chk(b, torch.float32)
    if layout == GEMM_NT:
        (M, K), Nn = a.shape, b.shape[0]
    elif layout == GEMM_NN:
        (M, K), Nn = a.shape, b.shape[1]
    else:
        (K, M), Nn = a.shape, b.shape[1]
    if out is None:
        out = torch.empty(M, Nn, dtype=torch.float32, device=a.device)
    epi = N.GemmEpilogue(None, None, None, alpha, 0, int(accumulate))
    rc = N.cuda().ainic_gemm_tf32(layout, a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0), out.data_ptr(), out.stride(0),
                                  M, Nn, K, C.byref(epi), block_n, max_ctas, N.stream_ptr())
    N.check(rc, "gemm_tf32")
    _count()
    return out


def mx_sf_bytes(rows: int, k: int) -> int:
    return N.cuda().ainic_mx_sf_bytes(rows, k)


def mx_quantize_rows(x: torch.Tensor, out: tuple | None = None):
    """x [rows, K] (fp32 / bf16, K contiguous) -> (q uint8 [rows, K] e4m3, scale atoms uint8): one ue8m0 scale per 32
    consecutive values of a row, stored in the tcgen05 scale-factor atom layout (gemm_kinds.cu).  ``out`` = a
    (q, scale atoms) pair from an earlier call with the same shape, to reuse its buffers."""
    assert x.is_cuda and x.dim() == 2 and x.stride(1) == 1 and x.dtype in (torch.float32, torch.bfloat16)
    rows, k = x.shape
    if out is not None:
        q, sf = out
        assert q.shape == (rows, k) and q.dtype == torch.uint8 and sf.numel() == mx_sf_bytes(rows, k)
    else:
        q = torch.empty(rows, k, dtype=torch.uint8, device=x.device)
        sf = torch.empty(mx_sf_bytes(rows, k), dtype=torch.uint8, device=x.device)
    N.check(N.cuda().ainic_mx_quantize_rows(x.data_ptr(), int(x.dtype == torch.bfloat16), rows, k, x.stride(0), q.data_ptr(),
                                            sf.data_ptr(), N.stream_ptr()), "mx_quantize_rows")
    _count(2)
    return q, sf


def mx_quantize_cols(x: torch.Tensor):
    """x [K, cols] (K is the slow dimension) -> (q uint8 [K, cols], scale atoms): one scale per 32 consecutive ROWS of a
    column - the operand form of the backward-weight product, where K is the minibatch."""
    assert x.is_cuda and x.dim() == 2 and x.stride(1) == 1 and x.dtype in (torch.float32, torch.bfloat16)
    k, cols = x.shape
    q = torch.empty(k, cols, dtype=torch.uint8, device=x.device)
    sf = torch.empty(mx_sf_bytes(cols, k), dtype=torch.uint8, device=x.device)
    N.check(N.cuda().ainic_mx_quantize_cols(x.data_ptr(), int(x.dtype == torch.bfloat16), k, cols, x.stride(0), q.data_ptr(),
                                            sf.data_ptr(), N.stream_ptr()), "mx_quantize_cols")
    _count(2)
    return q, sf


def gemm_mxfp8(layout: int, a: torch.Tensor, sfa: torch.Tensor, b: torch.Tensor, sfb: torch.Tensor, out_dtype=torch.bfloat16,
               out: torch.Tensor | None = None, block_n: int = 0, max_ctas: int = 0) -> torch.Tensor:
    """e4m3 operands + ue8m0 block scales, scaled inside the tensor core (tcgen05.mma.kind::mxf8f6f4.block_scale)."""
    _chk(a, torch.uint8)
    _chk(b, torch.uint8)
    if layout == GEMM_NT:
        (M, K), Nn = a.shape, b.shape[0]
    elif layout == GEMM_NN:
        (M, K), Nn = a.shape, b.shape[1]
    else:
        (K, M), Nn = a.shape, b.shape[1]
    if out is None:
        out = torch.empty(M, Nn, dtype=out_dtype, device=a.device)
    rc = N.cuda().ainic_gemm_mxfp8(layout, a.data_ptr(), a.stride(0), sfa.data_ptr(), b.data_ptr(), b.stride(0), sfb.data_ptr(),
                                   out.data_ptr(), out.stride(0), int(out.dtype == torch.float32), M, Nn, K, None, block_n,
                                   max_ctas, N.stream_ptr())
    N.check(rc, "gemm_mxfp8")
    _count()
    return out
