"""Property tests of the Torch7 file codec (utils/t7.py): arbitrary nested tables of numbers, strings,
booleans and tensors of every storage type survive binary and ASCII round trips; shared tensors stay
shared (the format memoises objects); truncated or corrupted input fails with an exception, never a hang
or a silent partial result."""
import math

import pytest
import torch
from hypothesis import given, settings, strategies as st

from tadnn.utils import t7

DTYPES = [torch.float32, torch.float64, torch.int64, torch.int32, torch.int16, torch.uint8, torch.int8]


@st.composite
def tensors(draw):
    dtype = draw(st.sampled_from(DTYPES))
    shape = draw(st.lists(st.integers(1, 4), min_size=1, max_size=3))
    n = math.prod(shape)
    if dtype.is_floating_point:
        vals = draw(st.lists(st.floats(-1e3, 1e3, allow_nan=False, width=32), min_size=n, max_size=n))
    else:
        lo, hi = (0, 255) if dtype == torch.uint8 else (-100, 100)
        vals = draw(st.lists(st.integers(lo, hi), min_size=n, max_size=n))
    t = torch.tensor(vals, dtype=dtype).reshape(shape)
    if len(shape) == 2 and draw(st.booleans()):
        t = t.t()  # non-contiguous view: strides must be written, not assumed
    return t


scalars = st.one_of(st.integers(-2**31, 2**31), st.floats(-1e6, 1e6, allow_nan=False), st.booleans(),
                    st.text(alphabet=st.characters(min_codepoint=32, max_codepoint=126), max_size=12))
values = st.recursive(st.one_of(scalars, tensors()),
                      lambda kids: st.dictionaries(st.one_of(st.integers(1, 20), st.text("abcdefgh", min_size=1, max_size=5)), kids, max_size=4),
                      max_leaves=10)


def _same(a, b):
    if torch.is_tensor(a):
        return torch.is_tensor(b) and a.dtype == b.dtype and a.shape == b.shape and torch.equal(a, b)
    if isinstance(a, dict):
        return isinstance(b, dict) and set(a) == set(b) and all(_same(a[k], b[k]) for k in a)
    if isinstance(a, bool) or isinstance(b, bool):
        return a is b
    if isinstance(a, (int, float)):
        return isinstance(b, (int, float)) and float(a) == float(b)  # Lua has one number type
    return a == b


@settings(max_examples=150, deadline=None, derandomize=True)
@given(obj=values, mode=st.sampled_from(["binary", "ascii"]))
def test_roundtrip_of_arbitrary_nested_values(obj, mode):
    back = t7.loads(t7.dumps(obj, mode), mode)
    assert _same(obj, back), (obj, back)


@pytest.mark.parametrize("mode", ["binary", "ascii"])
def test_shared_tensor_is_written_once_and_stays_shared(mode):
    w = torch.arange(6, dtype=torch.float32).reshape(2, 3)
    blob = t7.dumps({"a": w, "b": w, "view": w[1]}, mode)
    assert len(blob) < len(t7.dumps({"a": w, "b": w.clone(), "view": w[1].clone()}, mode))  # memoised, not duplicated
    back = t7.loads(blob, mode)
    back["a"][0, 0] = 42.0
    assert back["b"][0, 0] == 42.0           # same tensor object
    back["a"][1, 2] = -7.0
    assert back["view"][2] == -7.0           # a view of the same storage


@settings(max_examples=150, deadline=None, derandomize=True)
@given(obj=values, cut=st.floats(0.0, 0.98))
def test_truncated_binary_input_raises(obj, cut):
    blob = t7.dumps(obj, "binary")
    k = int(len(blob) * cut)
    if k >= len(blob):
        return
    try:
        back = t7.loads(blob[:k], "binary")
    except Exception:  # noqa: BLE001 - any exception is acceptable; a partial object is not
        return
    # a prefix can itself be a complete value only for scalars whose tail bytes were padding-free
    assert not isinstance(obj, dict) and not torch.is_tensor(obj) or _same(obj, back)


@settings(max_examples=150, deadline=None, derandomize=True)
@given(junk=st.binary(min_size=0, max_size=64))
def test_random_bytes_never_hang_or_crash_the_interpreter(junk):
    try:
        t7.loads(junk, "binary")
    except Exception:  # noqa: BLE001
        pass
    try:
        t7.loads(junk, "ascii")
    except Exception:  # noqa: BLE001
        pass


def test_zero_dim_and_empty_tensors():
    back = t7.loads(t7.dumps({"s": torch.tensor(2.5), "e": torch.empty(0)}))
    assert back["s"].shape == (1,) and back["s"].item() == 2.5  # no 0-dim tensors in Torch7
    assert back["e"].numel() == 0


def test_small_view_of_a_big_arena_gets_its_own_storage():
    arena = torch.zeros(1 << 20)
    v = arena[12345:12345 + 16]
    v.copy_(torch.arange(16.0))
    blob = t7.dumps({"v": v, "far": arena[-8:]})  # two small tensors 4 MB apart in one storage
    assert len(blob) < 4096  # the arena between them is not dragged into the file
    back = t7.loads(blob)
    assert torch.equal(back["v"], torch.arange(16.0)) and torch.equal(back["far"], torch.zeros(8))
