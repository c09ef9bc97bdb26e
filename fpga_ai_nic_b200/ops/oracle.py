"""CPU oracle wrappers (torch CPU tensors in, torch CPU tensors out).

The reference has no test suite; these restatements are the golden model:
bit-exact BFP codec (hw/bf16_to_bfp_core.sv, hw/bfp_to_bf16_core.sv, hw/bfp_adapter.sv), the ring
schedule of hw/all_reduce.sv:891-1086, and the MLP math of sw/mlp_mpi_example_f32.cpp:700-790.
"""
from __future__ import annotations

import torch

from .. import _native as N

HW_EXACT, CLEAN = 0, 1
CODEC_F32, CODEC_BFP, CODEC_MXFP8, CODEC_BF16 = 0, 1, 2, 3


def _f32(t: torch.Tensor) -> torch.Tensor:
    assert t.device.type == "cpu"
    return t.detach().to(torch.float32).contiguous()


def bfp_bytes(count: int) -> int:
    return N.oracle().ainic_ref_bfp_bytes(count)


def mx_bytes(count: int) -> int:
    return N.oracle().ainic_ref_mx_bytes(count)


def bfp_compress(x: torch.Tensor, mode: int = HW_EXACT) -> torch.Tensor:
    x = _f32(x).view(-1)
    out = torch.zeros(bfp_bytes(x.numel()), dtype=torch.uint8)
    N.oracle().ainic_ref_bfp_compress(x.data_ptr(), x.numel(), out.data_ptr(), mode)
    return out


def bfp_decompress(buf: torch.Tensor, count: int, mode: int = HW_EXACT) -> torch.Tensor:
    buf = buf.contiguous()
    out = torch.empty(count, dtype=torch.float32)
    N.oracle().ainic_ref_bfp_decompress(buf.data_ptr(), count, out.data_ptr(), mode)
    return out


def mx_compress(x: torch.Tensor) -> torch.Tensor:
    x = _f32(x).view(-1)
    out = torch.zeros(mx_bytes(x.numel()), dtype=torch.uint8)
    N.oracle().ainic_ref_mx_compress(x.data_ptr(), x.numel(), out.data_ptr())
    return out


def mx_decompress(buf: torch.Tensor, count: int) -> torch.Tensor:
    buf = buf.contiguous()
    out = torch.empty(count, dtype=torch.float32)
    N.oracle().ainic_ref_mx_decompress(buf.data_ptr(), count, out.data_ptr())
    return out


def bfp_roundtrip(x: torch.Tensor, mode: int = HW_EXACT) -> torch.Tensor:
    y = _f32(x).clone().view(-1)
    N.oracle().ainic_ref_bfp_roundtrip(y.data_ptr(), y.numel(), mode)
    return y.view(x.shape)


def allreduce(grads: torch.Tensor, algo: str = "exact", codec: int = CODEC_F32, mode: int = HW_EXACT) -> torch.Tensor:
    """grads: [nranks, count] -> what every rank applies, [nranks, count]."""
    g = _f32(grads)
    nranks, count = g.shape
    out = torch.empty_like(g)
    lib = N.oracle()
    if algo == "exact":
        lib.ainic_ref_allreduce_exact(g.data_ptr(), nranks, count, out.data_ptr())
    elif algo == "ring":
        assert codec in (CODEC_F32, CODEC_BFP)
        lib.ainic_ref_allreduce_ring(g.data_ptr(), nranks, count, int(codec == CODEC_BFP), mode, out.data_ptr())
    elif algo == "oneshot":
        lib.ainic_ref_allreduce_oneshot(g.data_ptr(), nranks, count, codec, mode, out.data_ptr())
    elif algo == "twoshot":
        lib.ainic_ref_allreduce_twoshot(g.data_ptr(), nranks, count, codec, mode, out.data_ptr())
    else:
        raise ValueError(algo)
    return out


def bucket_plan(layer_bytes, bucket_bytes: int, schedule=None):
    """The GPU trainer's gradient-bucket plan (csrc/api/bucket_plan.h), computed by the native code:
    [(lowest layer, highest layer), ...] in issue order (last layer first)."""
    import ctypes as C
    n = len(layer_bytes)
    lb = (C.c_size_t * n)(*[int(b) for b in layer_bytes])
    sched = (C.c_int * len(schedule))(*schedule) if schedule else None
    lo, hi = (C.c_int * n)(), (C.c_int * n)()
    N.oracle().ainic_ref_bucket_plan(n, lb, int(bucket_bytes), sched, len(schedule) if schedule else 0, lo, hi)
    return [(lo[l], hi[l]) for l in range(n - 1, -1, -1) if lo[l] == l]


def sgd(w: torch.Tensor, g: torch.Tensor, lr: float) -> torch.Tensor:
    w = _f32(w).clone()
    g = _f32(g)
    N.oracle().ainic_ref_sgd(w.data_ptr(), g.data_ptr(), w.numel(), lr)
    return w


def fc_fwd(x, w, bias=None, fuse: int = 0):
    x, w = _f32(x), _f32(w)
    mb, k = x.shape
    n = w.shape[0]
    y = torch.empty(mb, n)
    mask = torch.zeros(mb, n, dtype=torch.uint8)
    b = _f32(bias) if bias is not None else None
    N.oracle().ainic_ref_fc_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr() if b is not None else None, y.data_ptr(),
                                mask.data_ptr(), mb, k, n, fuse)
    return y, mask


def fc_bwd(x, w, dy, mask=None, fuse: int = 0, need_dx: bool = True):
    x, w, dy = _f32(x), _f32(w), _f32(dy).clone()
    mb, k = x.shape
    n = w.shape[0]
    dx = torch.empty(mb, k) if need_dx else None
    dw = torch.empty(n, k)
    db = torch.empty(n)
    N.oracle().ainic_ref_fc_bwd(x.data_ptr(), w.data_ptr(), dy.data_ptr(), mask.data_ptr() if mask is not None else None,
                                dx.data_ptr() if dx is not None else None, dw.data_ptr(), db.data_ptr(), mb, k, n, fuse)
    return dx, dw, db


def softmax_loss(logits, labels, loss_weight: float = 0.1):
    logits = _f32(logits)
    labels = labels.to(torch.int32).contiguous()
    mb, n = logits.shape
    probs = torch.empty(mb, n)
    dl = torch.empty(mb, n)
    loss = N.oracle().ainic_ref_softmax_loss(logits.data_ptr(), labels.data_ptr(), probs.data_ptr(), dl.data_ptr(), mb, n,
                                             loss_weight)
    return float(loss), probs, dl


def init_buf(n: int, seed: int = 1, init_pos: bool = False, init_one: bool = False) -> torch.Tensor:
    out = torch.empty(n, dtype=torch.float32)
    N.oracle().ainic_ref_init_buf(out.data_ptr(), n, seed, int(init_pos), int(init_one))
    return out
