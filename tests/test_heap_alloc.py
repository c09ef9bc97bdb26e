"""csrc/comm/heap_alloc.h (the symmetric heap's offset allocator) through the CPU hooks of the oracle library:
determinism (equal call sequences give equal offsets - what keeps the ranks' heaps symmetric), reuse after free,
coalescing, alignment, exhaustion, and a randomised model check against a brute-force overlap test."""
import ctypes as C
import random

import pytest

from fpga_ai_nic_b200 import _native as N

NONE = (1 << 64) - 1


@pytest.fixture()
def lib():
    return N.oracle()


class Heap:
    def __init__(self, lib, cap):
        self.lib, self.h = lib, lib.ainic_ref_heap_new(cap)

    def alloc(self, n, align=256):
        return self.lib.ainic_ref_heap_alloc(self.h, n, align)

    def free(self, off):
        return self.lib.ainic_ref_heap_free(self.h, off)

    def stats(self):
        out = (C.c_size_t * 3)()
        self.lib.ainic_ref_heap_stats(self.h, out)
        return {"in_use": out[0], "high_water": out[1], "free_blocks": out[2]}

    def __del__(self):
        self.lib.ainic_ref_heap_delete(self.h)


def test_bump_then_reuse_and_coalesce(lib):
    h = Heap(lib, 1 << 20)
    a = h.alloc(1000)            # rounded up to 1024
    b = h.alloc(4096, 4096)      # leaves the alignment gap [1024, 4096) on the free list
    c = h.alloc(300)             # first fit: goes into that gap
    assert a == 0 and b == 4096 and c == 1024
    assert h.stats() == {"in_use": 1024 + 4096 + 512, "high_water": 8192, "free_blocks": 1}
    e = h.alloc(8192)
    assert e == 8192
    assert h.free(b) == 1 and h.free(b) == 0                 # double free is refused
    assert h.stats()["free_blocks"] == 1                     # gap remainder + b merged into one block [1536, 8192)
    d = h.alloc(6656)                                        # exactly that block
    assert d == 1536 and h.stats()["free_blocks"] == 0
    assert h.free(e) == 1 and h.stats()["high_water"] == 8192    # a freed block at the end goes back to the tail
    assert h.free(d) == 1 and h.free(c) == 1 and h.free(a) == 1
    assert h.stats() == {"in_use": 0, "high_water": 0, "free_blocks": 0}


def test_exhaustion_and_alignment(lib):
    h = Heap(lib, 64 << 10)
    assert h.alloc(0) == NONE
    x = h.alloc(60 << 10, 4096)
    assert x == 0 and h.alloc(8 << 10) == NONE               # does not fit, state unchanged
    assert h.alloc(4 << 10) == 60 << 10
    assert h.alloc(1) == NONE
    h.free(x)
    y = h.alloc(256, 32 << 10)                               # aligned inside the freed block
    assert y % (32 << 10) == 0


def test_same_sequence_same_offsets_and_no_overlap(lib):
    """Two 'ranks' replay one random alloc / free sequence: identical offsets; live blocks never overlap."""
    rng = random.Random(7)
    ops = []
    live_count = 0
    for _ in range(400):
        if live_count and rng.random() < 0.45:
            ops.append(("free", rng.randrange(live_count)))
            live_count -= 1
        else:
            ops.append(("alloc", rng.choice([100, 256, 5000, 70000, 1 << 20]), rng.choice([256, 4096, 1 << 16])))
            live_count += 1
    traces = []
    for _rank in range(2):
        h = Heap(lib, 256 << 20)
        live, trace = [], []
        for op in ops:
            if op[0] == "alloc":
                off = h.alloc(op[1], op[2])
                assert off != NONE and off % op[2] == 0
                size = (op[1] + 255) // 256 * 256
                for (o, s) in live:
                    assert off + size <= o or o + s <= off, "live blocks overlap"
                live.append((off, size))
                trace.append(off)
            else:
                o, s = live.pop(op[1])
                assert h.free(o) == 1
        assert h.stats()["in_use"] == sum(s for _, s in live)
        traces.append(trace)
    assert traces[0] == traces[1]
