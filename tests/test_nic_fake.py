"""Command-interface semantics on the host fake device (no GPU): eight round-robin done slots,
in-order completion, <= 8 outstanding requests, wait with timeout, abort, per-call count, MPI-style
in-place sum.  Reference behaviour: sw/mlp_mpi_example_f32.cpp:114-180, hw/all_reduce.sv:1368-1375."""
import threading

import pytest
import torch

from fpga_ai_nic_b200.ops import oracle as O
from fpga_ai_nic_b200.parallel.nic import Nic, NicError


def _run_ranks(n, fn):
    errs = []

    def wrap(r):
        try:
            fn(r)
        except Exception as e:  # pragma: no cover
            errs.append((r, e))

    th = [threading.Thread(target=wrap, args=(r,)) for r in range(n)]
    [t.start() for t in th]
    [t.join(60) for t in th]
    assert not errs, errs


@pytest.mark.parametrize("nranks", [1, 2, 3])
@pytest.mark.parametrize("codec,algo", [("f32", "twoshot"), ("bfp", "twoshot"), ("bfp", "ring"), ("mxfp8", "oneshot")])
def test_fake_nic_allreduce_and_update(nranks, codec, algo):
    count = 8192 * 3 + 100
    g = torch.randn(nranks, count, generator=torch.Generator().manual_seed(0)) * 1e-3
    w0 = torch.randn(count, generator=torch.Generator().manual_seed(1))
    nics = [Nic("fake") for _ in range(nranks)]
    Nic.link(nics)
    bufs = [g[r].clone() for r in range(nranks)]
    ws = [w0.clone() for _ in range(nranks)]
    code = {"f32": O.CODEC_F32, "bfp": O.CODEC_BFP, "mxfp8": O.CODEC_MXFP8}[codec]
    want = O.allreduce(g, "exact" if codec == "f32" else algo, code, O.CLEAN)

    def rank_main(r):
        nics[r].setup(count, codec=codec, bfp_mode=O.CLEAN, algo=algo, lr=0.1)
        rq = nics[r].all_reduce(bufs[r], ws[r], flags=1)
        assert rq.idx == 0 and rq.count == count
        nics[r].wait(rq, 30)
        assert nics[r].done_word(0) == 1

    _run_ranks(nranks, rank_main)
    for r in range(nranks):
        assert torch.equal(ws[r], O.sgd(w0, want[r], 0.1))
        assert nics[r].latency_ns() > 0
    [n.close() for n in nics]


def test_done_slots_round_robin_in_order_and_busy():
    a, b = Nic("fake"), Nic("fake")
    Nic.link([a, b])
    a.setup(1024, codec="f32", algo="twoshot")
    b.setup(1024, codec="f32", algo="twoshot")
    bufs = [torch.ones(1024) * i for i in range(9)]
    rqs = [a.iallreduce(bufs[i]) for i in range(8)]          # peer never posts -> nothing completes
    assert [r.idx for r in rqs] == list(range(8))
    assert [r.seq for r in rqs] == list(range(8))
    with pytest.raises(NicError) as e:
        a.iallreduce(bufs[8])                                # ninth outstanding request is refused
    assert e.value.rc == -4
    with pytest.raises(NicError) as e:
        a.wait(rqs[0], timeout_s=0.2)                        # the reference would spin forever (D10)
    assert e.value.rc == -3
    # now the peer catches up; completions arrive in issue order
    pbufs = [torch.ones(1024) for _ in range(9)]            # requests are asynchronous: buffers must stay alive
    peer = [b.iallreduce(pbufs[i]) for i in range(8)]
    for i in range(8):
        a.wait(rqs[i], 30)
        b.wait(peer[i], 30)
        assert torch.equal(bufs[i], torch.ones(1024) * i + 1)
    r9 = a.iallreduce(bufs[8])                               # slot 0 is reused by sequence number 8
    assert r9.idx == 0 and r9.seq == 8
    pb = b.iallreduce(pbufs[8])
    a.wait(r9, 30)
    b.wait(pb, 30)
    a.close()
    b.close()


def test_abort_unblocks_wait():
    a, b = Nic("fake"), Nic("fake")
    Nic.link([a, b])
    a.setup(512, codec="f32")
    keep = torch.ones(512)
    rq = a.iallreduce(keep)
    t = threading.Timer(0.2, a.abort)
    t.start()
    with pytest.raises(NicError) as e:
        a.wait(rq, timeout_s=20)
    assert e.value.rc == -5
    with pytest.raises(NicError):
        a.iallreduce(torch.ones(512))
    a.close()
    b.close()


def test_per_call_count_differs_per_layer():
    """The reference programs one global length; here every request carries its own."""
    n = Nic("fake")
    n.setup(0, codec="f32", lr=0.5)
    for count in (512, 2048, 100):
        g, w = torch.ones(count), torch.zeros(count)
        rq = n.all_reduce(g, w)
        n.wait(rq, 10)
        assert torch.equal(w, torch.full((count,), -0.5))
    n.close()
