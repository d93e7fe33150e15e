"""The symmetric heap's allocator (parallel/symm.py) on the CPU: first fit, free, coalescing,
determinism.  No GPU and no native library involved — only the bookkeeping."""
import random

import pytest

from tadnn.parallel.symm import SymmContext


def _ctx(heap=1 << 20):
    c = SymmContext()
    c.heap_bytes = heap
    c._free = [(0, heap)]
    c._live = {}
    return c


def test_alloc_free_coalesces_back_to_one_block():
    c = _ctx()
    offs = [c.alloc(n) for n in (100, 4096, 1, 70000, 256)]
    assert len(set(offs)) == 5 and all(o % 256 == 0 for o in offs)
    assert c.live_bytes() + c.free_bytes() == c.heap_bytes
    for o in (offs[1], offs[3], offs[0], offs[4], offs[2]):
        c.free(o)
    assert c._free == [(0, c.heap_bytes)] and c.live_bytes() == 0


def test_cycling_twenty_batch_sizes_does_not_exhaust_the_heap():
    """What a ragged last batch or a sweep over B does to the layer plans: every new size allocates
    three buffers and releases the previous ones; a bump allocator died here."""
    c = _ctx(8 << 20)
    prev = []
    for B in list(range(100, 2100, 100)) * 3:
        cur = [c.alloc(B * 1024), c.alloc(B * 512), c.alloc(B * 64)]
        for o in prev:
            c.free(o)
        prev = cur
    for o in prev:
        c.free(o)
    assert c._free == [(0, c.heap_bytes)]


def test_exhaustion_and_bad_free_are_errors():
    c = _ctx(1 << 16)
    a = c.alloc(1 << 15)
    with pytest.raises(MemoryError, match="symmetric heap exhausted"):
        c.alloc((1 << 15) + 256)
    with pytest.raises(ValueError):
        c.free(a + 256)
    c.free(a)
    with pytest.raises(ValueError):
        c.free(a)  # double free
    assert c.alloc(1 << 16) == 0


def test_same_call_sequence_gives_same_offsets():
    rng = random.Random(7)
    ops = []
    live = 0
    for _ in range(300):
        if live and rng.random() < 0.45:
            ops.append(("free", rng.randrange(live)))
            live -= 1
        else:
            ops.append(("alloc", rng.randrange(1, 50000)))
            live += 1

    def replay():
        c, held, trace = _ctx(64 << 20), [], []
        for op, v in ops:
            if op == "alloc":
                held.append(c.alloc(v))
                trace.append(held[-1])
            else:
                c.free(held.pop(v))
        return trace

    assert replay() == replay()


def test_alignment_request_is_honoured():
    c = _ctx()
    c.alloc(300)
    o = c.alloc(1000, align=4096)
    assert o % 4096 == 0
