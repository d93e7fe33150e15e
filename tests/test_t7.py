""".t7 reader/writer: round trips + hand-built golden bytes."""
import struct

import numpy as np
import pytest
import torch

from tadnn import mpi, nn
from tadnn.utils import t7


def _i(*v):
    return struct.pack(f"<{len(v)}i", *v)


def _q(*v):
    return struct.pack(f"<{len(v)}q", *v)


def _s(s):
    return _i(len(s)) + s.encode()


def test_golden_binary_number_string_bool_nil():
    assert t7.loads(_i(1) + struct.pack("<d", 3.5)) == 3.5
    assert t7.loads(_i(1) + struct.pack("<d", 7.0)) == 7
    assert t7.loads(_i(2) + _s("hello")) == "hello"
    assert t7.loads(_i(5) + _i(1)) is True
    assert t7.loads(_i(0)) is None


def test_golden_binary_float_tensor():
    # torch.FloatTensor 2x3, contiguous, storage of 6 floats
    data = np.arange(6, dtype=np.float32)
    blob = (_i(4) + _i(1) + _s("V 1") + _s("torch.FloatTensor") + _i(2) + _q(2, 3) + _q(3, 1) + _q(1)
            + _i(4) + _i(2) + _s("V 1") + _s("torch.FloatStorage") + _q(6) + data.tobytes())
    t = t7.loads(blob)
    assert t.dtype == torch.float32 and t.shape == (2, 3)
    assert torch.equal(t, torch.arange(6.).reshape(2, 3))
    # our writer produces the identical byte stream
    assert t7.dumps(torch.arange(6.).reshape(2, 3)) == blob


def test_golden_binary_table_with_memo_and_class():
    # { weight = <tensor#2>, alias = <same tensor> } wrapped in nn.Linear class object
    data = np.array([1.0, 2.0], dtype=np.float64)
    tensor = (_i(4) + _i(3) + _s("V 1") + _s("torch.DoubleTensor") + _i(1) + _q(2) + _q(1) + _q(1)
              + _i(4) + _i(4) + _s("V 1") + _s("torch.DoubleStorage") + _q(2) + data.tobytes())
    blob = (_i(4) + _i(1) + _s("V 1") + _s("nn.Linear") + _i(3) + _i(2) + _i(2)
            + _i(2) + _s("weight") + tensor
            + _i(2) + _s("alias") + _i(4) + _i(3))
    o = t7.loads(blob)
    assert isinstance(o, t7.T7Object) and o.className == "nn.Linear"
    assert o["weight"] is o["alias"]
    assert o["weight"].tolist() == [1.0, 2.0]


def test_golden_ascii_cifar_like_table():
    txt = b"3\n1\n2\n2\n4\ndata\n4\n2\n3\nV 1\n16\ntorch.ByteTensor\n2\n2 3\n3 1\n1\n4\n3\n3\nV 1\n17\ntorch.ByteStorage\n6\n1 2 3 4 5 6\n2\n6\nlabels\n4\n4\n3\nV 1\n16\ntorch.ByteTensor\n1\n3\n1\n1\n4\n5\n3\nV 1\n17\ntorch.ByteStorage\n3\n7 8 9\n"
    o = t7.loads(txt, "ascii")
    assert o["data"].dtype == torch.uint8 and o["data"].tolist() == [[1, 2, 3], [4, 5, 6]]
    assert o["labels"].tolist() == [7, 8, 9]


def test_function_payload_is_refused():
    with pytest.raises(ValueError):
        t7.loads(_i(6) + _i(1))


@pytest.mark.parametrize("mode", ["binary", "ascii"])
def test_roundtrip_objects(mode):
    obj = {"a": 1, "b": [1.5, "x", True, None], "t": torch.randn(3, 4), "l": torch.arange(5),
           "nc": torch.randn(4, 6).t()}
    back = t7.loads(t7.dumps(obj, mode), mode)
    assert back["a"] == 1 and t7.table_to_list(back["b"])[:3] == [1.5, "x", True]
    assert torch.allclose(back["t"], obj["t"]) and torch.equal(back["l"], obj["l"])
    assert torch.allclose(back["nc"], obj["nc"])


@pytest.mark.parametrize("mode", ["binary", "ascii"])
def test_model_roundtrip_dense_and_cnn(mode, tmp_path):
    m = (nn.Sequential().add(nn.SpatialConvolutionMM(3, 4, 5, 5)).add(nn.ReLU()).add(nn.SpatialMaxPooling(3, 3, 3, 3))
         .add(nn.View(4 * 9 * 9)).add(nn.Dropout(0.5)).add(nn.Linear(324, 10)).add(nn.LogSoftMax()))
    p = str(tmp_path / "m.t7")
    t7.save_model(p, m, mode)
    m2 = t7.load_model(p, mode)
    m.eval(), m2.eval()
    x = torch.randn(2, 3, 32, 32)
    assert torch.allclose(m(x), m2(x), atol=1e-6)
    o = t7.load(p, mode)
    assert [c.className for c in t7.table_to_list(o["modules"])] == [
        "nn.SpatialConvolutionMM", "nn.ReLU", "nn.SpatialMaxPooling", "nn.View", "nn.Dropout", "nn.Linear", "nn.LogSoftMax"]
    assert o["modules"][1]["weight"].shape == (4, 75)  # Torch7 SpatialConvolutionMM layout


def test_mp_model_roundtrip_per_rank_shards(tmp_path):
    world = mpi.FakeWorld(2)

    def body(h):
        torch.manual_seed(3 + h.rank())
        m = (nn.Sequential().add(nn.MPInitialReshape(16)).add(nn.MPInitialLinear(16, 8)).add(nn.MPTanh())
             .add(nn.MPBaseLinear(8, 4)))
        p = str(tmp_path / f"mp_rank{h.rank()}.t7")
        t7.save_model(p, m)
        o = t7.load(p)
        names = [c.className for c in t7.table_to_list(o["modules"])]
        assert names == ["nn.MPInitialReshape", "nn.MPInitialLinear", "nn.MPTanh", "nn.MPBaseLinear"]
        assert o["modules"][2]["weight"].shape == (8, 8) and o["modules"][4]["weight"].shape == (4, 4)
        m2 = t7.load_model(p)
        x = torch.randn(3, 4, 4, generator=torch.Generator().manual_seed(0))
        return (m(x) - m2(x)).abs().max().item()

    assert max(world.run(body)) < 1e-6
