"""The allocation-free all-reduce emulation (``ainic_allreduce_emulate_noalloc``).

Some callers of this library cannot allocate device memory at the moment they need an all-reduce:
a server that gave all of HBM to its own pool at start-up (a fresh cudaMalloc simply fails there), and a
loop that must not trigger cudaFree's device-wide synchronisation.  For them the library offers one entry point whose whole scratch area is a fixed
2 MiB ``__device__`` array instead of per-call cudaMalloc/cudaFree.

The test drives that entry point through the C ABI (ctypes) exactly as such a caller would and
checks, twice in a row, that it produces bit-identical results to the allocating entry point
``ainic_allreduce_emulate``.  The second call is the interesting one: it finds the previous call's
flags, counters and staging bytes still sitting in the shared scratch array.  The file name makes
pytest collect it last, so a ``-x`` run has already exercised every other test when it gets here;
it must also pass on its own in a fresh process, where its first call is the one that resolves the
scratch address on the device.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_RANKS, MAX_COUNT = 4, 32768   # AINIC_NOALLOC_MAX_RANKS / AINIC_NOALLOC_MAX_COUNT in include/ainic.h


def _lib():
    from fpga_ai_nic_b200 import _native
    return _native.cuda()


def _call_noalloc(lib, grads, weights, shadow, gout, lr, codec, mode, algo, ctas=3):
    nranks, count = grads.shape
    err = C.create_string_buffer(256)
    rc = lib.ainic_allreduce_emulate_noalloc(nranks, grads.data_ptr(), weights.data_ptr(), shadow.data_ptr(),
                                             gout.data_ptr(), count, lr, codec, mode, algo, ctas,
                                             torch.cuda.current_stream().cuda_stream, err, len(err))
    return rc, err.value.decode()


def _call_alloc(lib, grads, weights, shadow, gout, lr, codec, mode, algo, ctas=3):
    nranks, count = grads.shape
    return lib.ainic_allreduce_emulate(nranks, grads.data_ptr(), weights.data_ptr(), shadow.data_ptr(), gout.data_ptr(),
                                       count, lr, codec, mode, algo, ctas, torch.cuda.current_stream().cuda_stream)


# (algorithm, codec, bfp mode): 1 oneshot / 2 twoshot / 3 ring; 0 f32 / 1 bfp / 2 mxfp8 / 3 bf16
CASES = [(2, 1, 1), (1, 1, 0), (3, 1, 0), (2, 0, 0), (3, 0, 0), (1, 2, 0), (2, 3, 0)]


@pytest.mark.parametrize("algo,codec,mode", CASES)
@pytest.mark.parametrize("nranks,count", [(MAX_RANKS, MAX_COUNT), (3, 8192 * 3 - 76), (2, 700)])
def test_noalloc_matches_allocating_entry_point_twice(algo, codec, mode, nranks, count):
    lib = _lib()
    gen = torch.Generator().manual_seed(1000 * algo + 10 * codec + nranks)
    for call in range(2):                      # two calls in a row share the static scratch
        g = torch.randn(nranks, count, generator=gen) * 1e-3
        w = torch.randn(nranks, count, generator=gen)

        def run(fn):
            gd, wd = g.cuda(), w.cuda()
            sh = torch.zeros(nranks, count, dtype=torch.bfloat16, device="cuda")
            go = torch.zeros(nranks, count, device="cuda")
            out = fn(lib, gd, wd, sh, go, 0.1, codec, mode, algo)
            torch.cuda.synchronize()
            return out, go.cpu(), wd.cpu(), sh.cpu()

        rc_ref, go_ref, w_ref, sh_ref = run(_call_alloc)
        assert rc_ref == 0
        (rc, msg), go, wn, sh = run(_call_noalloc)
        assert rc == 0, msg
        # same tolerance as the other all-reduce tests: bit-for-bit
        assert torch.equal(go.view(torch.int32), go_ref.view(torch.int32)), (call, (go - go_ref).abs().max())
        assert torch.equal(wn, w_ref), call
        assert torch.equal(sh, sh_ref), call


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    lib = _lib()
    g = torch.zeros(MAX_RANKS + 1, 512, device="cuda")
    w, go = torch.zeros_like(g), torch.zeros_like(g)
    sh = torch.zeros(MAX_RANKS + 1, 512, dtype=torch.bfloat16, device="cuda")
    rc, msg = _call_noalloc(lib, g, w, sh, go, 0.1, 1, 1, 2)
    assert rc == -1 and "at most 4" in msg
    g = torch.zeros(2, MAX_COUNT + 512, device="cuda")
    w, go = torch.zeros_like(g), torch.zeros_like(g)
    sh = torch.zeros(2, MAX_COUNT + 512, dtype=torch.bfloat16, device="cuda")
    rc, msg = _call_noalloc(lib, g, w, sh, go, 0.1, 1, 1, 2)
    assert rc == -1 and "at most 32768" in msg


@pytest.mark.parametrize("algo_name,algo,codec,mode", [("twoshot", 2, 1, 1), ("ring", 3, 1, 0), ("oneshot", 1, 2, 0)])
def test_noalloc_matches_cpu_oracle_directly(algo_name, algo, codec, mode):
    """Not only 'same as the allocating entry point': the bytes coming out of the static-scratch path
    are compared with the CPU golden model itself (gradient applied by every rank, SGD result)."""
    from fpga_ai_nic_b200.ops import oracle as O
    lib = _lib()
    nranks, count = 3, 8192 * 3 - 76
    gen = torch.Generator().manual_seed(4242 + algo)
    g = torch.randn(nranks, count, generator=gen) * 1e-3
    w0 = (torch.randn(count, generator=gen) * 0.05).repeat(nranks, 1)
    want = O.allreduce(g, algo_name, codec, mode)
    gd, wd = g.cuda(), w0.clone().cuda()
    sh = torch.zeros(nranks, count, dtype=torch.bfloat16, device="cuda")
    go = torch.zeros(nranks, count, device="cuda")
    rc, msg = _call_noalloc(lib, gd, wd, sh, go, 0.1, codec, mode, algo)
    torch.cuda.synchronize()
    assert rc == 0, msg
    assert torch.equal(go.cpu().view(torch.int32), want.view(torch.int32))
    for r in range(nranks):
        assert torch.equal(wd[r].cpu(), O.sgd(w0[r], want[r], 0.1))
