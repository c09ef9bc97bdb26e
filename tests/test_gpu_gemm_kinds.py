"""tf32 and MX block-scaled tcgen05 GEMMs (csrc/gemm/gemm_kinds.cu) against plain fp32 PyTorch.

tf32: the operands keep 10 mantissa bits inside the tensor core, so the reference is an fp32 matmul and
the tolerance is tf32's.  MX: the quantiser's output is decoded on the host (e4m3 value x 2^(scale-127),
scale found through the atom layout) and multiplied in fp32 - the GEMM must reproduce that product up to
fp32 summation order, which proves operand layout, scale placement in tensor memory and the sf_id
selection per K-block.  The quantiser itself is compared bit for bit with the MXFP8 wire codec."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from fpga_ai_nic_b200.ops import kernels
    return kernels


def _sf_lookup(sf, rows, k):
    """scale byte of (row, k-block) -> tensor [rows, k/32] from the atom-tiled array."""
    r = torch.arange(rows).view(-1, 1)
    kb = torch.arange(k // 32).view(1, -1)
    off = ((r // 128) * (k // 128) + kb // 4) * 512 + (r % 32) * 16 + ((r % 128) // 32) * 4 + kb % 4
    return sf.cpu()[off.flatten()].view(rows, k // 32)


def _dequant_rows(q, sf):
    rows, k = q.shape
    s = _sf_lookup(sf, rows, k).to(torch.int32)
    scale = torch.where(s > 0, torch.pow(2.0, (s - 127).float()), torch.zeros(()))
    v = q.cpu().view(torch.float8_e4m3fn).float()
    return v * scale.repeat_interleave(32, dim=1)


def _dequant_cols(q, sf):
    return _dequant_rows(q.t().contiguous(), sf).t().contiguous()


@pytest.mark.parametrize("layout", ["NT", "NN", "TN"])
@pytest.mark.parametrize("shape", [(256, 256, 128), (1792, 2048, 2048), (384, 200, 136), (132, 72, 36)])
@pytest.mark.parametrize("block_n", [0, 128])
def test_tf32_gemm_matches_fp32(K, layout, shape, block_n):
    M, Nn, Kd = shape
    g = torch.Generator().manual_seed(M + Nn)
    lay = {"NT": K.GEMM_NT, "NN": K.GEMM_NN, "TN": K.GEMM_TN}[layout]
    a = torch.randn((M, Kd) if layout != "TN" else (Kd, M), generator=g).cuda()
    b = torch.randn((Nn, Kd) if layout == "NT" else (Kd, Nn), generator=g).cuda()
    out = K.gemm_tf32(lay, a, b, block_n=block_n)
    A = a if layout != "TN" else a.t()
    B = b.t() if layout == "NT" else b
    want = (A.double() @ B.double()).float()
    torch.cuda.synchronize()
    err = (out - want).abs().max().item()
    assert err <= 4e-3 * want.abs().max().item() + 1e-3 * Kd ** 0.5, err
    # alpha / accumulate epilogue
    out2 = K.gemm_tf32(lay, a, b, out=out.clone(), alpha=0.5, accumulate=True, block_n=block_n)
    assert (out2 - 1.5 * out).abs().max().item() <= 1e-4 * out.abs().max().item() + 1e-5


def test_mx_quantiser_matches_wire_codec(K):
    from fpga_ai_nic_b200.ops import oracle as O
    g = torch.Generator().manual_seed(5)
    x = (torch.randn(256, 1024, generator=g) * torch.logspace(-3, 2, 1024).view(1, -1))
    x[3, 64:96] = 0.0
    q, sf = K.mx_quantize_rows(x.cuda())
    frames = O.mx_compress(x.flatten()).view(-1, 528)          # 512 e4m3 bytes + 16 scale bytes per 512 values
    assert torch.equal(q.cpu().flatten(), frames[:, :512].flatten())
    assert torch.equal(_sf_lookup(sf, 256, 1024).flatten(), frames[:, 512:].flatten())
    # column-wise blocks: the transpose of the row-wise result
    qc, sfc = K.mx_quantize_cols(x.t().contiguous().cuda())
    assert torch.equal(qc.cpu().t().contiguous(), q.cpu())
    assert torch.equal(_sf_lookup(sfc, 256, 1024), _sf_lookup(sf, 256, 1024))


@pytest.mark.parametrize("layout", ["NT", "NN", "TN"])
@pytest.mark.parametrize("shape", [(128, 128, 128), (256, 384, 512), (1792, 2048, 2048), (208, 272, 256)])
@pytest.mark.parametrize("block_n", [128, 256])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_mx_block_scaled_gemm(K, layout, shape, block_n, out_dtype):
    M, Nn, Kd = shape
    g = torch.Generator().manual_seed(7 * M + Kd)
    lay = {"NT": K.GEMM_NT, "NN": K.GEMM_NN, "TN": K.GEMM_TN}[layout]
    # per-block dynamic range, so that a misplaced scale factor cannot hide
    amp_m = torch.logspace(-2, 1, Kd // 32).repeat_interleave(32)
    A = torch.randn(M, Kd, generator=g) * amp_m
    B = torch.randn(Nn, Kd, generator=g) * amp_m.flip(0) * (1 + torch.arange(Nn).view(-1, 1) % 7)
    if layout == "TN":
        a, sfa = K.mx_quantize_cols(A.t().contiguous().cuda())
        Ad = _dequant_cols(a, sfa).t()
    else:
        a, sfa = K.mx_quantize_rows(A.cuda())
        Ad = _dequant_rows(a, sfa)
    if layout == "NT":
        b, sfb = K.mx_quantize_rows(B.cuda())
        Bd = _dequant_rows(b, sfb)
    else:
        b, sfb = K.mx_quantize_cols(B.t().contiguous().cuda())
        Bd = _dequant_cols(b, sfb).t()
    out = K.gemm_mxfp8(lay, a, sfa, b, sfb, out_dtype=out_dtype, block_n=block_n)
    torch.cuda.synchronize()
    want = (Ad.double() @ Bd.double().t()).float()
    tol = 2e-5 if out_dtype == torch.float32 else 1e-2
    err = (out.float().cpu() - want).abs().max().item()
    assert err <= tol * want.abs().max().item(), (err, want.abs().max().item())
    # and the quantised product is a sane approximation of the exact one
    exact = A @ B.t()
    assert ((want - exact).norm() / exact.norm()).item() < 0.06
