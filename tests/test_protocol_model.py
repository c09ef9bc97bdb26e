"""Interleaving model of the cross-GPU flag protocol of csrc/comm/allreduce.cuh (DESIGN.md section 3,
"Cross-GPU protocol").  compute-sanitizer is closed on this GPU pool, so the buffer-reuse argument is
checked here instead: every (rank, CTA) is a coroutine that performs the kernel's memory operations in
program order, a random scheduler interleaves them, and every read asserts that it sees the data of
the CURRENT call - never a leftover of an older call and never something a faster rank has already
overwritten for a later call.  The negative controls drop one ingredient of the protocol each and
must be caught, which shows the model can see the races it is looking for.

What the reference has in this place: dual-clock FIFOs with overflow / underflow checking and credit
based back-pressure (hw/fifo.v:87-88, hw/all_reduce.sv:410-427) - hardware flow control instead of a
software protocol."""
import random

import pytest


class Violation(Exception):
    pass


class World:
    def __init__(self, n, ctas):
        self.n, self.ctas = n, ctas
        # flags[dst][phase][src][cta] = last epoch signalled by (src, cta) to dst
        self.flags = [[[[0] * ctas for _ in range(n)] for _ in range(3)] for _ in range(n)]
        self.stage_c = [[[0] * ctas for _ in range(2)] for _ in range(n)]   # [rank][parity][cta] = epoch of the data
        self.stage_d = [[0] * ctas for _ in range(n)]                      # [rank][cta]


def twoshot(w, r, b, calls, start_barrier=True, parity=True):
    """One CTA of the compressed two-shot kernel, `calls` consecutive launches (twoshot_body)."""
    for epoch in range(1, calls + 1):
        par = (epoch & 1) if parity else 0
        w.stage_c[r][par][b] = epoch                       # phase 0: compress my gradient into my stage C
        yield
        for p in range(w.n):                               # signal_all(phase 0)
            if p != r:
                w.flags[p][0][r][b] = epoch
                yield
        if start_barrier:
            while any(w.flags[r][0][p][b] < epoch for p in range(w.n) if p != r):   # wait_all(phase 0)
                yield
        for p in range(w.n):                               # phase 1: reduce my shard from every rank's stage C
            if w.stage_c[p][par][b] != epoch:
                raise Violation(f"rank {r} cta {b} call {epoch}: stage C of rank {p} holds call {w.stage_c[p][par][b]}")
            yield
        w.stage_d[r][b] = epoch                            # publish the reduced shard
        yield
        for p in range(w.n):                               # signal_all / wait_all(phase 1)
            if p != r:
                w.flags[p][1][r][b] = epoch
                yield
        while any(w.flags[r][1][p][b] < epoch for p in range(w.n) if p != r):
            yield
        for p in range(w.n):                               # phase 2: gather every other rank's shard
            if p != r:
                if w.stage_d[p][b] != epoch:
                    raise Violation(f"rank {r} cta {b} call {epoch}: stage D of rank {p} holds call {w.stage_d[p][b]}")
                yield


def oneshot_precompressed(w, r, b, calls, double_buffer=True):
    """One CTA of the one-shot kernel reading gradients the GEMM epilogue compressed (precompressed == 1):
    the producer writes buffer step & 1 before the launch, there is NO closing barrier."""
    for epoch in range(1, calls + 1):
        par = (epoch & 1) if double_buffer else 0
        w.stage_c[r][par][b] = epoch                       # the weight-gradient GEMM of this step
        yield
        for p in range(w.n):
            if p != r:
                w.flags[p][0][r][b] = epoch
                yield
        while any(w.flags[r][0][p][b] < epoch for p in range(w.n) if p != r):
            yield
        for p in range(w.n):
            if w.stage_c[p][par][b] != epoch:
                raise Violation(f"rank {r} cta {b} call {epoch}: frames of rank {p} belong to call {w.stage_c[p][par][b]}")
            yield


def twoshot_scattered(w, r, b, calls, double_buffer=True, start_barrier=True):
    """GEMM -> reduce-scatter epilogue + two-shot (precompressed == 2): before the launch this rank's GEMM
    stores its frames of every shard into the OWNER's inbox (slot [my rank]); the kernel's first flag
    exchange tells the owners that the stores are complete; the reduce phase reads the local inbox."""
    if not hasattr(w, "inbox"):
        w.inbox = [[[[0] * w.ctas for _ in range(w.n)] for _ in range(2)] for _ in range(w.n)]   # [owner][parity][src][cta]
    for epoch in range(1, calls + 1):
        par = (epoch & 1) if double_buffer else 0
        for p in range(w.n):                               # weight-gradient GEMM epilogue: NVLink stores
            w.inbox[p][par][r][b] = epoch
            yield
        for p in range(w.n):
            if p != r:
                w.flags[p][0][r][b] = epoch
                yield
        if start_barrier:
            while any(w.flags[r][0][p][b] < epoch for p in range(w.n) if p != r):
                yield
        for p in range(w.n):                               # reduce my shard: local reads only
            if w.inbox[r][par][p][b] != epoch:
                raise Violation(f"rank {r} cta {b} call {epoch}: inbox slot of rank {p} holds call {w.inbox[r][par][p][b]}")
            yield
        w.stage_d[r][b] = epoch
        yield
        for p in range(w.n):
            if p != r:
                w.flags[p][1][r][b] = epoch
                yield
        while any(w.flags[r][1][p][b] < epoch for p in range(w.n) if p != r):
            yield
        for p in range(w.n):
            if p != r:
                if w.stage_d[p][b] != epoch:
                    raise Violation(f"rank {r} cta {b} call {epoch}: stage D of rank {p} holds call {w.stage_d[p][b]}")
                yield


def run(body, n, ctas, calls, seed, **kw):
    rng = random.Random(seed)
    w = World(n, ctas)
    live = [body(w, r, b, calls, **kw) for r in range(n) for b in range(ctas)]
    # a biased scheduler: some coroutines are "slow GPUs" and get picked rarely, which is what exposes reuse races
    weight = [rng.choice((1, 1, 5, 25)) for _ in live]
    steps = 0
    while live:
        i = rng.choices(range(len(live)), weights=weight)[0]
        try:
            next(live[i])
        except StopIteration:
            live.pop(i)
            weight.pop(i)
        steps += 1
        assert steps < 2_000_000, "deadlock in the model"
    return steps


@pytest.mark.parametrize("n", [2, 3, 8])
def test_twoshot_protocol_has_no_reuse_race(n):
    for seed in range(40):
        run(twoshot, n, 2, 6, seed)


@pytest.mark.parametrize("n", [2, 4])
def test_precompressed_oneshot_is_safe_with_two_buffers(n):
    for seed in range(40):
        run(oneshot_precompressed, n, 2, 6, seed)


@pytest.mark.parametrize("n", [2, 4, 8])
@pytest.mark.parametrize("double_buffer", [True, False])
def test_scattered_twoshot_is_safe(n, double_buffer):
    """Safe even with ONE inbox as long as a rank's next GEMM runs after its own previous call on that
    bucket (the trainer's `updated` dependency): the second flag exchange orders it after every owner's
    reads.  The two inboxes per bucket are margin for callers that relax that dependency."""
    for seed in range(30):
        run(twoshot_scattered, n, 2, 6, seed, double_buffer=double_buffer)


def _caught(body, **kw):
    hits = 0
    for seed in range(60):
        try:
            run(body, 3, 2, 6, seed, **kw)
        except Violation:
            hits += 1
    return hits


def test_negative_controls_are_caught():
    # without the start-of-call flag exchange a fast rank reads a slow rank's stage C before it is written
    assert _caught(twoshot, start_barrier=False) > 0
    # with a single GEMM output buffer and no closing barrier a fast rank's next step overwrites frames a
    # slow rank is still reading
    assert _caught(oneshot_precompressed, double_buffer=False) > 0
    # scattered input: without the first exchange an owner reduces slots its peers have not stored yet
    assert _caught(twoshot_scattered, start_barrier=False) > 0


def test_twoshot_alone_would_not_need_the_parity_halves():
    """Documentation of the argument in DESIGN.md: the second flag exchange of two-shot already orders a
    rank's next compression after every peer's reads; the two halves of stage C exist for one-shot
    (and the GEMM-written buffers), which has no closing exchange - see the negative control above."""
    for seed in range(40):
        run(twoshot, 3, 2, 6, seed, parity=False)


# ------------------------------------------------------------------------------------------ flow all-reduce
# csrc/comm/flow.cuh: ONE persistent kernel per step.  Roles inside a rank run concurrently: the GEMMs (producer)
# write chunk k's compressed frames and ring its doorbell; the forwarder turns doorbells into "readable" flags on
# every rank; reduce warps wait for every rank's flag of chunk k, reduce their shard, publish it into stage D and
# the LAST of them (ticket) posts "published"; gather warps wait for the peers' "published" flags and read the
# peers' stage D.  Kernel boundaries are stream order: a rank's step e + 1 starts when all of its roles of step e
# have finished.
class FlowWorld:
    def __init__(self, n, chunks, reducers, gatherers):
        self.n, self.chunks, self.nr, self.ng = n, chunks, reducers, gatherers
        self.cg = [[0] * chunks for _ in range(n)]              # [rank][chunk] = step whose frames the buffer holds
        self.doorbell = [[0] * chunks for _ in range(n)]
        self.ready = [[[0] * n for _ in range(chunks)] for _ in range(n)]     # [dst][chunk][src]
        self.pub = [[[0] * n for _ in range(chunks)] for _ in range(n)]
        self.stage_d = [[0] * chunks for _ in range(n)]
        self.ticket = [[0] * chunks for _ in range(n)]
        self.done_roles = [0] * n                               # roles of the current step that have finished
        self.step = [1] * n


def _flow_role(w, r, steps, body, nroles):
    """Run `body(step)` once per step; a rank's roles advance to the next step together (kernel boundary)."""
    for e in range(1, steps + 1):
        while w.step[r] != e:
            yield
        yield from body(e)
        w.done_roles[r] += 1
        if w.done_roles[r] == nroles:
            w.done_roles[r] = 0
            w.step[r] = e + 1
        yield


def flow_coroutines(w, steps, wait_doorbell=True, wait_pub=True, last_warp_posts=True):
    n, C = w.n, w.chunks
    nroles = 2 + w.nr + w.ng
    cos = []
    for r in range(n):
        def producer(e, r=r):
            for k in range(C):
                w.cg[r][k] = e
                yield
                w.doorbell[r][k] = e
                yield

        def forwarder(e, r=r):
            for k in range(C):
                while wait_doorbell and w.doorbell[r][k] < e:
                    yield
                for p in range(n):
                    w.ready[p][k][r] = e
                    yield

        def reducer(e, r=r):
            for k in range(C):
                while any(w.ready[r][k][p] < e for p in range(n)):
                    yield
                for p in range(n):
                    if w.cg[p][k] != e:
                        raise Violation(f"rank {r} step {e} chunk {k}: frames of rank {p} belong to step {w.cg[p][k]}")
                    yield
                w.stage_d[r][k] = e          # (every reduce warp writes its own frames of the shard; one cell stands for them)
                yield
                w.ticket[r][k] += 1
                last = w.ticket[r][k] == w.nr
                if last:
                    w.ticket[r][k] = 0
                if last or not last_warp_posts:
                    for p in range(n):
                        if p != r:
                            w.pub[p][k][r] = e
                            yield

        def gatherer(e, r=r):
            for k in range(C):
                while wait_pub and any(w.pub[r][k][p] < e for p in range(n) if p != r):
                    yield
                for p in range(n):
                    if p != r:
                        if w.stage_d[p][k] != e:
                            raise Violation(f"rank {r} step {e} chunk {k}: stage D of rank {p} holds step {w.stage_d[p][k]}")
                        yield

        cos.append(_flow_role(w, r, steps, producer, nroles))
        cos.append(_flow_role(w, r, steps, forwarder, nroles))
        cos += [_flow_role(w, r, steps, reducer, nroles) for _ in range(w.nr)]
        cos += [_flow_role(w, r, steps, gatherer, nroles) for _ in range(w.ng)]
    return cos


def run_flow_model(n, seed, steps=4, chunks=3, reducers=2, gatherers=2, **kw):
    rng = random.Random(seed)
    w = FlowWorld(n, chunks, reducers, gatherers)
    live = flow_coroutines(w, steps, **kw)
    weight = [rng.choice((1, 1, 5, 25)) for _ in live]
    budget = 0
    while live:
        i = rng.choices(range(len(live)), weights=weight)[0]
        try:
            next(live[i])
        except StopIteration:
            live.pop(i)
            weight.pop(i)
        budget += 1
        assert budget < 3_000_000, "deadlock in the model"


@pytest.mark.parametrize("n", [2, 3, 4])
def test_flow_protocol_has_no_race_and_no_deadlock(n):
    for seed in range(25):
        run_flow_model(n, seed)


def test_flow_negative_controls_are_caught():
    def caught(**kw):
        hits = 0
        for seed in range(40):
            try:
                run_flow_model(3, seed, **kw)
            except Violation:
                hits += 1
        return hits
    # a forwarder that does not wait for the GEMMs' doorbell announces frames that are not written yet
    assert caught(wait_doorbell=False) > 0
    # gather warps that do not wait for "published" read the previous step's shard
    assert caught(wait_pub=False) > 0
    # if every reduce warp posted "published" instead of only the last one (the ticket), a peer could finish the step
    # while slower reduce warps of this rank are still reading that peer's frames - and its next step's GEMMs
    # overwrite them (the model keeps ONE frame buffer per rank; the real trainer alternates two, which is margin)
    assert caught(last_warp_posts=False) > 0
