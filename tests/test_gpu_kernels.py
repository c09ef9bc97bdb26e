"""GPU numerics: every hand-written kernel against the CPU oracle / a plain PyTorch fp32 reference."""
import pytest
import torch

from fpga_ai_nic_b200.ops import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    from fpga_ai_nic_b200.ops import kernels
    return kernels


def _grad(n, seed=0, scale=1e-3):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, generator=g) * scale
    x[::97] = 0.0
    return x


@pytest.mark.parametrize("count", [512, 8192, 512 * 37 + 100, 1 << 20])
@pytest.mark.parametrize("mode", [O.HW_EXACT, O.CLEAN])
def test_bfp_codec_bit_exact(K, count, mode):
    x = _grad(count)
    x[:16] = 0.0
    want = O.bfp_compress(x, mode)
    got = K.bfp_compress(x.cuda(), mode).cpu()
    # bytes of the padded tail are defined (zeros encode deterministically)
    assert torch.equal(got, want)
    dec = K.bfp_decompress(want.cuda(), count, mode).cpu()
    assert torch.equal(dec.view(torch.int32), O.bfp_decompress(want, count, mode).view(torch.int32))


@pytest.mark.parametrize("count", [512, 512 * 9 + 7, 1 << 18])
def test_mx_codec_bit_exact(K, count):
    x = _grad(count, seed=1)
    want = O.mx_compress(x)
    got = K.mx_compress(x.cuda()).cpu()
    assert torch.equal(got, want)
    dec = K.mx_decompress(want.cuda(), count).cpu()
    assert torch.equal(dec, O.mx_decompress(want, count))


@pytest.mark.parametrize("count", [4, 1023, 1 << 20])
def test_sgd_matches_fmaf(K, count):
    w, g = _grad(count, 2, 0.05), _grad(count, 3)
    want = O.sgd(w, g, 0.1)
    wd, sh = w.cuda(), torch.empty(count, dtype=torch.bfloat16, device="cuda")
    K.sgd_(wd, g.cuda(), 0.1, sh)
    assert torch.equal(wd.cpu(), want)
    assert torch.equal(sh.cpu(), want.to(torch.bfloat16))


def _ref_mm(a, b):
    return a.float() @ b.float()


GEMM_SHAPES = [(128, 256, 64), (256, 512, 256), (384, 1024, 1088), (1792, 2048, 2048), (200, 328, 72)]   # last: ragged in every dimension


@pytest.mark.parametrize("mb,n,k", GEMM_SHAPES)
@pytest.mark.parametrize("block_n", [0, 128, 208, 256, -128, -208, -256])   # < 0: 2-CTA kernel
def test_gemm_fwd(K, mb, n, k, block_n):
    torch.manual_seed(0)
    x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
    y = K.gemm_fwd(x, w, block_n=block_n)
    ref = _ref_mm(x, w.t())
    torch.cuda.synchronize()
    err = (y.float() - ref).abs().max().item()
    assert err <= 0.02 * ref.abs().max().item() + 1e-3, err


@pytest.mark.parametrize("mb,n,k", GEMM_SHAPES)
@pytest.mark.parametrize("block_n", [0, 128, 192, 256, -128, -256])
def test_gemm_bwd_data(K, mb, n, k, block_n):
    torch.manual_seed(1)
    dy = (torch.randn(mb, n, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
    dx = K.gemm_bwd_data(dy, w, block_n=block_n)
    ref = _ref_mm(dy, w)
    torch.cuda.synchronize()
    err = (dx.float() - ref).abs().max().item()
    assert err <= 0.02 * ref.abs().max().item() + 1e-3, err


@pytest.mark.parametrize("mb,n,k", GEMM_SHAPES)
@pytest.mark.parametrize("block_n", [0, 128, 192, 256, -128, -256])
def test_gemm_bwd_weight(K, mb, n, k, block_n):
    torch.manual_seed(2)
    dy = (torch.randn(mb, n, device="cuda") * 0.05).bfloat16()
    x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
    dw = K.gemm_bwd_weight(dy, x, block_n=block_n)
    ref = _ref_mm(dy.t(), x)
    torch.cuda.synchronize()
    err = (dw - ref).abs().max().item()
    assert err <= 2e-3 * ref.abs().max().item() + 1e-4, err
    # accumulate + alpha epilogue
    dw2 = dw.clone()
    K.gemm_bwd_weight(dy, x, out=dw2, accumulate=True, alpha=0.5, block_n=block_n)
    torch.cuda.synchronize()
    assert (dw2 - 1.5 * ref).abs().max().item() <= 4e-3 * ref.abs().max().item() + 1e-4


@pytest.mark.parametrize("codec,mode", [("bfp", O.CLEAN), ("bfp", O.HW_EXACT), ("mxfp8", 0)])
@pytest.mark.parametrize("mb,n,k,block_n", [(256, 512, 256, 0), (1792, 2048, 2048, 0), (384, 256, 512, 128), (384, 384, 1024, 192),
                                                 (256, 320, 1024, 256), (128, 2048, 4096, 256)])   # partial row tile; more tiles than SMs (direct + TMA epilogues mixed)
def test_gemm_bwd_weight_compressed_epilogue(K, codec, mode, mb, n, k, block_n):
    """The weight-gradient GEMM can emit the wire format directly: its bytes must equal the standalone
    codec applied to the fp32 gradient the same kernel writes (hw/bfp_adapter.sv behind the adder tree)."""
    torch.manual_seed(6)
    dy = (torch.randn(mb, n, device="cuda") * 0.05).bfloat16()
    x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
    dw = K.gemm_bwd_weight(dy, x, block_n=block_n)
    count = n * k
    if codec == "bfp":
        want = K.bfp_compress(dw.view(-1), mode)
        got = torch.zeros(K.bfp_bytes(count), dtype=torch.uint8, device="cuda")
        K.gemm_bwd_weight(dy, x, block_n=block_n, compressed_out=got, codec=K.CODEC_BFP, bfp_mode=mode)
    else:
        want = K.mx_compress(dw.view(-1))
        got = torch.zeros(K.mx_bytes(count), dtype=torch.uint8, device="cuda")
        K.gemm_bwd_weight(dy, x, block_n=block_n, compressed_out=got, codec=K.CODEC_MXFP8)
    torch.cuda.synchronize()
    assert torch.equal(got, want)


@pytest.mark.parametrize("codec", ["bfp", "mxfp8"])
@pytest.mark.parametrize("nranks,rank,frame0,extra", [(4, 2, 5, 3), (8, 7, 0, 0), (2, 0, 130, 1), (1, 0, 4, 0)])
def test_gemm_scatter_epilogue(K, codec, nranks, rank, frame0, extra):
    """GEMM -> reduce-scatter: every frame lands in its owner's inbox at slot [rank][frame % shard]
    (ainic_gemm_scatter_t); here the "peers" are nranks buffers on one GPU."""
    torch.manual_seed(8)
    mb, n, k = 256, 256, 512
    dy = (torch.randn(mb, n, device="cuda") * 0.05).bfloat16()
    x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
    frames = n * k // 512
    fb = 544 if codec == "bfp" else 528
    code = K.CODEC_BFP if codec == "bfp" else K.CODEC_MXFP8
    want = torch.zeros(frames * fb, dtype=torch.uint8, device="cuda")
    K.gemm_bwd_weight(dy, x, compressed_out=want, codec=code)
    total = frame0 + frames + extra                       # frames of the whole bucket
    shard = (total + nranks - 1) // nranks
    inbox = [torch.full((nranks * shard * fb,), 0xEE, dtype=torch.uint8, device="cuda") for _ in range(nranks)]
    K.gemm_bwd_weight(dy, x, codec=code, scatter=(inbox, rank, shard, frame0))
    torch.cuda.synchronize()
    want = want.view(frames, fb).cpu()
    got = [b.view(nranks, shard, fb).cpu() for b in inbox]
    touched = 0
    for f in range(frames):
        o, j = divmod(frame0 + f, shard)
        assert torch.equal(got[o][rank, j], want[f]), (f, o, j)
        touched += 1
    # nothing else was written
    assert sum(int((g != 0xEE).any(dim=-1).sum()) for g in got) <= touched


def test_gemm_fused_epilogues(K):
    torch.manual_seed(3)
    mb, n, k = 256, 512, 384
    x = (torch.randn(mb, k, device="cuda") * 0.5).bfloat16()
    w = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
    bias = torch.randn(n, device="cuda") * 0.1
    mask = torch.zeros(mb, n, dtype=torch.uint8, device="cuda")
    y = K.gemm_fwd(x, w, bias=bias, relu=True, mask_out=mask)
    pre = _ref_mm(x, w.t()) + bias
    ref = torch.relu(pre)
    torch.cuda.synchronize()
    assert (y.float() - ref).abs().max().item() <= 0.02 * ref.abs().max().item() + 1e-3
    # the mask may differ only where the pre-activation is within rounding of zero
    diff = (mask.bool() != (pre > 0)) & (pre.abs() > 1e-2)
    assert not bool(diff.any())
    dy = (torch.randn(mb, n, device="cuda") * 0.5).bfloat16()
    w2 = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
    mask_in = (torch.rand(mb, k, device="cuda") > 0.5).to(torch.uint8)
    dx = K.gemm_bwd_data(dy, w2, mask_in=mask_in)
    ref = _ref_mm(dy, w2) * mask_in.float()
    torch.cuda.synchronize()
    assert (dx.float() - ref).abs().max().item() <= 0.02 * ref.abs().max().item() + 1e-3


@pytest.mark.parametrize("mb,n", [(64, 2048), (1792, 2048), (33, 1000)])
def test_softmax_xent(K, mb, n):
    torch.manual_seed(4)
    logits = (torch.randn(mb, n) * 2).bfloat16()
    labels = torch.randint(0, n, (mb,), dtype=torch.int32)
    loss, probs, dl = O.softmax_loss(logits.float(), labels, 0.1)
    ls, dlg, pg = K.softmax_xent(logits.cuda(), labels.cuda(), 0.1, want_probs=True)
    torch.cuda.synchronize()
    assert abs(ls.item() / mb - loss) <= 2e-3 * abs(loss) + 1e-4
    assert (pg.float().cpu() - probs).abs().max().item() <= 1e-2 * probs.max().item()
    assert (dlg.float().cpu() - dl).abs().max().item() <= 1e-2 * dl.abs().max().item()


def test_bias_grad(K):
    torch.manual_seed(5)
    dy = (torch.randn(1792, 2048, device="cuda") * 0.1).bfloat16()
    got = K.bias_grad(dy)
    ref = dy.float().sum(0)
    torch.cuda.synchronize()
    assert (got - ref).abs().max().item() <= 1e-3 * ref.abs().max().item() + 1e-4
