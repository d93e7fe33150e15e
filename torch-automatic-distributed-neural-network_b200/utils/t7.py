"""Torch7 ``.t7`` reader/writer (binary and ASCII) and ``tadnn.nn`` <-> Torch7 ``nn`` model
conversion.

The reference itself only *loads* ``.t7`` files — ``torch.load(path, 'ascii')`` of the CIFAR
batches (cifar_example/sgd-torchad_nn-cifar.lua:146,152) — but its five layer classes are
registered with ``torch.class('nn.MPInitialLinear', 'nn.Linear')`` etc.
(nodemodule.lua:34,86,132,186,238), which is what makes models serialisable under those class
names.  This module speaks the Torch7 ``File`` object format so that models move between the
two worlds in the reference's per-rank parameter layout (``weight [o, i/P]``, ``bias [o]``).

Format (``torch/File.lua``): ``object := int32 type, payload``; types 0 nil, 1 number (f64),
2 string (int32 len + bytes), 3 table, 4 torch object, 5 boolean; tables and torch objects
carry an ``int32`` memo index and are written once.  A torch object is ``"V 1"``, class name,
then the class payload: tensors ``int32 nDim, int64 size[], int64 stride[], int64 offset
(1-based), storage object``; storages ``int64 n, raw elements``; Lua-defined classes (all of
``nn.*``) a table of their fields.  ASCII mode is the same stream as whitespace-separated
decimal text.
"""
from __future__ import annotations

import io
import struct
import warnings
from typing import Any, Dict, Optional

import numpy as np
import torch

TYPE_NIL, TYPE_NUMBER, TYPE_STRING, TYPE_TABLE, TYPE_TORCH, TYPE_BOOLEAN = 0, 1, 2, 3, 4, 5
TYPE_FUNCTION, TYPE_RECUR_FUNCTION, LEGACY_TYPE_RECUR_FUNCTION = 6, 8, 7

_TENSOR_DTYPES = {
    "Double": np.float64, "Float": np.float32, "Half": np.float16, "Long": np.int64,
    "Int": np.int32, "Short": np.int16, "Char": np.int8, "Byte": np.uint8,
    "Cuda": np.float32, "CudaDouble": np.float64, "CudaHalf": np.float16, "CudaLong": np.int64,
    "CudaInt": np.int32, "CudaByte": np.uint8,
}
_NP_TO_T7 = {np.dtype(np.float64): "Double", np.dtype(np.float32): "Float", np.dtype(np.float16): "Half",
             np.dtype(np.int64): "Long", np.dtype(np.int32): "Int", np.dtype(np.int16): "Short",
             np.dtype(np.int8): "Char", np.dtype(np.uint8): "Byte"}


class T7Object:
    """A Lua-defined torch class instance (e.g. ``nn.Linear``): class name + field table."""

    def __init__(self, className: str, fields: Optional[Dict[Any, Any]] = None):
        self.className = className
        self.fields: Dict[Any, Any] = fields if fields is not None else {}

    def __getitem__(self, k):
        return self.fields[k]

    def get(self, k, default=None):
        return self.fields.get(k, default)

    def __repr__(self):
        return f"T7Object({self.className}, keys={list(self.fields)})"


class T7Storage:
    """A ``torch.*Storage`` (e.g. the ``LongStorage`` in ``nn.Reshape.size``)."""

    def __init__(self, array, kind: Optional[str] = None):
        self.array = np.asarray(array)
        self.kind = kind or _NP_TO_T7[self.array.dtype]

    def tolist(self):
        return self.array.tolist()

    def __repr__(self):
        return f"T7Storage({self.kind}, {self.array.tolist()})"


def table_to_list(t: Dict[Any, Any]) -> list:
    """1-based array table -> Python list."""
    n = len(t)
    return [t[i] if i in t else t[float(i)] for i in range(1, n + 1)]


# ----------------------------------------------------------------------------
# reader
# ----------------------------------------------------------------------------
class _Reader:
    def __init__(self, data: bytes, ascii_mode: bool, long_size: int = 8):
        self.b = data
        self.pos = 0
        self.ascii = ascii_mode
        self.long_size = long_size
        self.memo: Dict[int, Any] = {}

    # -- primitives ------------------------------------------------------------
    def _token(self) -> bytes:
        b, n = self.b, len(self.b)
        i = self.pos
        while i < n and b[i] in b" \t\r\n":
            i += 1
        j = i
        while j < n and b[j] not in b" \t\r\n":
            j += 1
        if i == j:
            raise EOFError("unexpected end of .t7 stream")
        self.pos = j
        return b[i:j]

    def ints(self, n: int) -> list:
        if self.ascii:
            return [int(self._token()) for _ in range(n)]
        out = list(struct.unpack_from(f"<{n}i", self.b, self.pos))
        self.pos += 4 * n
        return out

    def longs(self, n: int) -> list:
        if self.ascii:
            return [int(self._token()) for _ in range(n)]
        fmt = "q" if self.long_size == 8 else "i"
        out = list(struct.unpack_from(f"<{n}{fmt}", self.b, self.pos))
        self.pos += self.long_size * n
        return out

    def double(self) -> float:
        if self.ascii:
            return float(self._token())
        (v,) = struct.unpack_from("<d", self.b, self.pos)
        self.pos += 8
        return v

    def string(self) -> str:
        (n,) = self.ints(1)
        if self.ascii:
            if self.pos < len(self.b) and self.b[self.pos:self.pos + 1] in (b"\n", b" "):
                self.pos += 1
        if n < 0 or self.pos + n > len(self.b):
            raise EOFError(f".t7 stream ends inside a {n}-byte string at byte {self.pos}")
        s = self.b[self.pos:self.pos + n]
        self.pos += n
        return s.decode("latin-1")

    def array(self, n: int, np_dtype) -> np.ndarray:
        dt = np.dtype(np_dtype)
        if self.ascii:
            if n >= 64:
                # Torch7 writes an array as one line of space-separated numbers: parse the line in C
                # (a CIFAR batch is 30.7 M tokens; the per-token loop below would take a minute)
                b, i = self.b, self.pos
                while i < len(b) and b[i] in b" \t\r\n":
                    i += 1
                nl = b.find(b"\n", i)
                nl = len(b) if nl < 0 else nl
                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")
                    try:
                        a = np.fromstring(b[i:nl], dtype=np.float64 if dt.kind == "f" else np.int64, sep=" ")
                    except Exception:  # noqa: BLE001 - any parser problem: take the token loop
                        a = None
                if a is not None and a.size == n:
                    self.pos = nl
                    return a.astype(dt)
            if dt.kind == "f":
                return np.array([float(self._token()) for _ in range(n)], dtype=dt)
            return np.array([int(self._token()) for _ in range(n)], dtype=dt)
        a = np.frombuffer(self.b, dtype=dt.newbyteorder("<"), count=n, offset=self.pos).copy()
        self.pos += n * dt.itemsize
        return a

    # -- objects ---------------------------------------------------------------
    def obj(self):
        (t,) = self.ints(1)
        if t == TYPE_NIL:
            return None
        if t == TYPE_NUMBER:
            v = self.double()
            return int(v) if float(v).is_integer() and abs(v) < 2 ** 53 else v
        if t == TYPE_STRING:
            return self.string()
        if t == TYPE_BOOLEAN:
            return bool(self.ints(1)[0])
        if t == TYPE_TABLE:
            (idx,) = self.ints(1)
            if idx in self.memo:
                return self.memo[idx]
            tab: Dict[Any, Any] = {}
            self.memo[idx] = tab
            (n,) = self.ints(1)
            for _ in range(n):
                k = self.obj()
                v = self.obj()
                tab[k] = v
            return tab
        if t == TYPE_TORCH:
            (idx,) = self.ints(1)
            if idx in self.memo:
                return self.memo[idx]
            version = self.string()
            if version.startswith("V "):
                cls = self.string()
            else:  # pre-versioning files: the string just read IS the class name
                cls = version
            return self._torch_object(idx, cls)
        if t in (TYPE_FUNCTION, TYPE_RECUR_FUNCTION, LEGACY_TYPE_RECUR_FUNCTION):
            raise ValueError(".t7 stream contains a serialised Lua function; refusing to load it")
        raise ValueError(f"unknown .t7 object type {t} at byte {self.pos}")

    def _torch_object(self, idx: int, cls: str):
        if cls.startswith("torch.") and cls.endswith("Tensor"):
            kind = cls[len("torch."):-len("Tensor")]
            (nd,) = self.ints(1)
            size = self.longs(nd)
            stride = self.longs(nd)
            (off,) = self.longs(1)
            storage = self.obj()
            if storage is None or nd == 0:
                t = torch.empty(0, dtype=torch.from_numpy(np.zeros(0, _TENSOR_DTYPES[kind])).dtype)
            else:
                base = torch.from_numpy(storage.array)
                t = torch.as_strided(base, size, stride, off - 1)
            self.memo[idx] = t
            return t
        if cls.startswith("torch.") and cls.endswith("Storage"):
            kind = cls[len("torch."):-len("Storage")]
            (n,) = self.longs(1)
            st = T7Storage(self.array(n, _TENSOR_DTYPES[kind]), kind)
            self.memo[idx] = st
            return st
        o = T7Object(cls)
        self.memo[idx] = o
        fields = self.obj()
        o.fields = fields if isinstance(fields, dict) else {"__payload": fields}
        return o


def loads(data: bytes, mode: str = "binary", long_size: int = 8):
    return _Reader(data, mode == "ascii", long_size).obj()


def load(path: str, mode: str = "binary", long_size: int = 8):
    """``torch.load(path, mode)``; ``mode`` is "binary" or "ascii"."""
    with open(path, "rb") as f:
        return loads(f.read(), mode, long_size)


# ----------------------------------------------------------------------------
# writer
# ----------------------------------------------------------------------------
class _Writer:
    def __init__(self, ascii_mode: bool):
        self.out = io.BytesIO()
        self.ascii = ascii_mode
        self.memo: Dict[int, int] = {}
        self.keep = []  # keep ids alive
        self.next = 1
        # storage sharing (Torch7 memoises storages, so views of one buffer -- e.g. everything behind
        # getParameters() -- stay views after a save/load): key -> [lo, hi, referenced elements], then
        # key -> T7Storage of that window once it has been materialised
        self.windows: Dict[Any, list] = {}
        self.storages: Dict[Any, "T7Storage"] = {}

    @staticmethod
    def _skey(t: torch.Tensor):
        return (t.untyped_storage().data_ptr(), t.dtype, str(t.device))

    @staticmethod
    def _extent(t: torch.Tensor):
        lo = t.storage_offset()
        return lo, lo + 1 + sum((n - 1) * st for n, st in zip(t.shape, t.stride()))

    def plan(self, o, seen=None):
        """Pre-pass over the object graph: the element window of every underlying storage that the
        tensors to be written actually cover."""
        seen = set() if seen is None else seen
        if id(o) in seen:
            return
        seen.add(id(o))
        if isinstance(o, torch.Tensor):
            if o.numel() > 0:
                lo, hi = self._extent(o)
                w = self.windows.setdefault(self._skey(o), [lo, hi, 0])
                w[0], w[1], w[2] = min(w[0], lo), max(w[1], hi), w[2] + o.numel()
        elif isinstance(o, dict):
            for k, v in o.items():
                self.plan(k, seen)
                self.plan(v, seen)
        elif isinstance(o, (list, tuple)) and not isinstance(o, torch.Size):
            for v in o:
                self.plan(v, seen)
        elif isinstance(o, T7Object):
            self.plan(o.fields, seen)

    def ints(self, vals):
        if self.ascii:
            self.out.write((" ".join(str(int(v)) for v in vals) + "\n").encode())
        else:
            self.out.write(struct.pack(f"<{len(vals)}i", *[int(v) for v in vals]))

    def longs(self, vals):
        if self.ascii:
            self.out.write((" ".join(str(int(v)) for v in vals) + "\n").encode())
        else:
            self.out.write(struct.pack(f"<{len(vals)}q", *[int(v) for v in vals]))

    def double(self, v: float):
        if self.ascii:
            self.out.write((repr(float(v)) + "\n").encode())
        else:
            self.out.write(struct.pack("<d", float(v)))

    def string(self, s: str):
        b = s.encode("latin-1")
        self.ints([len(b)])
        self.out.write(b)
        if self.ascii:
            self.out.write(b"\n")

    def array(self, a: np.ndarray):
        if self.ascii:
            if a.dtype.kind == "f":
                self.out.write((" ".join(repr(float(v)) for v in a.tolist()) + "\n").encode())
            else:
                self.out.write((" ".join(str(int(v)) for v in a.tolist()) + "\n").encode())
        else:
            self.out.write(np.ascontiguousarray(a).astype(a.dtype.newbyteorder("<"), copy=False).tobytes())

    def _index(self, o) -> Optional[int]:
        k = id(o)
        if k in self.memo:
            self.ints([self.memo[k]])
            return None
        self.memo[k] = self.next
        self.keep.append(o)
        self.ints([self.next])
        self.next += 1
        return self.memo[k]

    def obj(self, o):
        if o is None:
            self.ints([TYPE_NIL])
        elif isinstance(o, bool):
            self.ints([TYPE_BOOLEAN])
            self.ints([1 if o else 0])
        elif isinstance(o, (int, float)) and not isinstance(o, bool):
            self.ints([TYPE_NUMBER])
            self.double(float(o))
        elif isinstance(o, str):
            self.ints([TYPE_STRING])
            self.string(o)
        elif isinstance(o, (list, tuple)):
            self._table({i + 1: v for i, v in enumerate(o)}, o)
        elif isinstance(o, dict):
            self._table(o, o)
        elif isinstance(o, torch.Tensor):
            self._tensor(o)
        elif isinstance(o, T7Storage):
            self.ints([TYPE_TORCH])
            if self._index(o) is not None:
                self.string("V 1")
                self.string(f"torch.{o.kind}Storage")
                self.longs([o.array.size])
                self.array(o.array.astype(_TENSOR_DTYPES[o.kind]))
        elif isinstance(o, T7Object):
            self.ints([TYPE_TORCH])
            if self._index(o) is not None:
                self.string("V 1")
                self.string(o.className)
                self.obj(o.fields)
        elif isinstance(o, torch.Size):
            self.obj(T7Storage(np.array(list(o), dtype=np.int64), "Long"))
        else:
            raise TypeError(f"cannot serialise {type(o)} to .t7")

    def _table(self, d: dict, ident):
        self.ints([TYPE_TABLE])
        if self._index(ident) is None:
            return
        items = [(k, v) for k, v in d.items()]
        self.ints([len(items)])
        for k, v in items:
            self.obj(k)
            self.obj(v)

    def _tensor(self, t: torch.Tensor):
        self.ints([TYPE_TORCH])
        if self._index(t) is None:
            return
        src = t.detach()
        if src.dim() == 0:
            src = src.reshape(1)  # Torch7 has no 0-dimensional tensors: a scalar becomes a 1-element vector
        nd = src.dim() if src.numel() > 0 else 0
        key = self._skey(src) if nd > 0 else None
        win = self.windows.get(key) if nd > 0 else None
        # share the underlying storage unless the window is mostly data nobody references (a small
        # tensor carved out of a big arena): then the tensor gets a compact storage of its own
        share = win is not None and (win[1] - win[0]) <= 4 * win[2] + 1024
        if nd > 0 and share:
            st = self.storages.get(key)
            if st is None:
                flat = torch.as_strided(src, (win[1] - win[0],), (1,), win[0]).cpu()
                if flat.dtype == torch.bfloat16:
                    flat = flat.float()
                a = flat.contiguous().numpy()
                st = self.storages[key] = T7Storage(a, _NP_TO_T7[a.dtype])
            kind, shape, strides, off = st.kind, list(src.shape), list(src.stride()), src.storage_offset() - win[0] + 1
        else:
            c = src.cpu()
            if c.dtype == torch.bfloat16:
                c = c.float()
            a = c.contiguous().numpy()
            kind = _NP_TO_T7[a.dtype]
            shape, strides, off = list(a.shape), [int(x // a.itemsize) for x in a.strides], 1
            st = T7Storage(a.reshape(-1), kind) if nd > 0 else None
        self.string("V 1")
        self.string(f"torch.{kind}Tensor")
        self.ints([nd])
        if nd == 0:
            self.longs([1])  # storageOffset
            self.obj(None)
            return
        self.longs(shape)
        self.longs(strides)
        self.longs([off])
        self.obj(st)


def dumps(obj, mode: str = "binary") -> bytes:
    w = _Writer(mode == "ascii")
    w.plan(obj)
    w.obj(obj)
    return w.out.getvalue()


def save(path: str, obj, mode: str = "binary") -> None:
    """``torch.save(path, obj, mode)``."""
    with open(path, "wb") as f:
        f.write(dumps(obj, mode))


# ----------------------------------------------------------------------------
# tadnn.nn  <->  Torch7 nn
# ----------------------------------------------------------------------------
def _empty():
    return torch.empty(0)


def _module_fields(**kw):
    f = {"_type": "torch.FloatTensor", "output": _empty(), "gradInput": _empty(), "train": True}
    f.update(kw)
    return f


def model_to_t7(module: torch.nn.Module) -> T7Object:
    """Torch7 object tree of a ``tadnn.nn`` model, with the reference's class names and the
    inherited field tables (``weight``, ``bias``, ``gradWeight``, ``gradBias``, …)."""
    from .. import nn as tnn
    from ..nn import mp

    def grad_of(p):
        return p.grad.detach().clone() if p.grad is not None else torch.zeros_like(p)

    if isinstance(module, torch.nn.Sequential):
        mods = {i + 1: model_to_t7(m) for i, m in enumerate(module.children())}
        return T7Object("nn.Sequential", _module_fields(modules=mods))
    if isinstance(module, (mp._MPLinear, torch.nn.Linear)):
        name = getattr(module, "t7_name", "nn.Linear")
        f = _module_fields(weight=module.weight.detach().float().clone(), gradWeight=grad_of(module.weight).float())
        if module.bias is not None:
            f["bias"] = module.bias.detach().float().clone()
            f["gradBias"] = grad_of(module.bias).float()
        return T7Object(name, f)
    if isinstance(module, (mp.MPTanh, torch.nn.Tanh)):
        return T7Object(getattr(module, "t7_name", "nn.Tanh"), _module_fields())
    if isinstance(module, torch.nn.ReLU):
        return T7Object("nn.ReLU", _module_fields(threshold=0, val=0, inplace=False))
    if isinstance(module, (mp._MPReshape, tnn.Reshape)):
        size = T7Storage(np.array(module.size, dtype=np.int64), "Long")
        bs = T7Storage(np.array((1,) + tuple(module.size), dtype=np.int64), "Long")
        f = _module_fields(size=size, batchsize=bs, nelement=module.nelement)
        if module.batchMode is not None:
            f["batchMode"] = bool(module.batchMode)
        return T7Object(getattr(module, "t7_name", "nn.Reshape"), f)
    if isinstance(module, tnn.View):
        return T7Object("nn.View", _module_fields(size=T7Storage(np.array(module.size, dtype=np.int64), "Long"),
                                                  numElements=module.nelement))
    if isinstance(module, tnn.SpatialConvolutionMM):
        return T7Object("nn.SpatialConvolutionMM", _module_fields(
            nInputPlane=module.nInputPlane, nOutputPlane=module.nOutputPlane, kW=module.kW, kH=module.kH,
            dW=module.dW, dH=module.dH, padW=module.padW, padH=module.padH,
            weight=module.weight.detach().float().clone(), bias=module.bias.detach().float().clone(),
            gradWeight=grad_of(module.weight).float(), gradBias=grad_of(module.bias).float()))
    if isinstance(module, tnn.SpatialMaxPooling):
        return T7Object("nn.SpatialMaxPooling", _module_fields(
            kW=module.kW, kH=module.kH, dW=module.dW, dH=module.dH, padW=module.padW, padH=module.padH,
            ceil_mode=False))
    if isinstance(module, torch.nn.Dropout):
        return T7Object("nn.Dropout", _module_fields(p=module.p, v2=True, noise=_empty()))
    if isinstance(module, tnn.LogSoftMax):
        return T7Object("nn.LogSoftMax", _module_fields())
    raise TypeError(f"no Torch7 counterpart registered for {type(module).__name__}")


def model_from_t7(o: T7Object, mpi=None) -> torch.nn.Module:
    """Rebuild a ``tadnn.nn`` model from a Torch7 object tree.  ``nn.MP*Linear`` objects carry
    per-rank shards ``weight [o, i/P]``; the module is created for the *current* world size and
    must match (``i/P`` is checked)."""
    from .. import mpi as _mpi
    from .. import nn as tnn
    h = mpi or _mpi.current()
    c = o.className

    def num(k, d=None):
        v = o.get(k, d)
        return int(v) if v is not None else d

    if c == "nn.Sequential":
        seq = tnn.Sequential()
        for m in table_to_list(o["modules"]):
            seq.add(model_from_t7(m, mpi))
        return seq
    if c in ("nn.Linear", "nn.MPInitialLinear", "nn.MPBaseLinear"):
        w = o["weight"].float()
        if c == "nn.Linear":
            m = tnn.Linear(w.shape[1], w.shape[0], bias=o.get("bias") is not None)
        else:
            cls = tnn.MPInitialLinear if c == "nn.MPInitialLinear" else tnn.MPBaseLinear
            m = cls(w.shape[1] * h.size(), w.shape[0], mpi=mpi)
            assert m.shard == w.shape[1], f"{c}: shard width {w.shape[1]} != i/P {m.shard}"
        with torch.no_grad():
            m.weight.copy_(w)
            if o.get("bias") is not None:
                m.bias.copy_(o["bias"].float())
        return m
    if c == "nn.Tanh":
        return tnn.Tanh()
    if c == "nn.MPTanh":
        return tnn.MPTanh(mpi=mpi)
    if c == "nn.ReLU":
        return tnn.ReLU()
    if c in ("nn.Reshape", "nn.MPInitialReshape", "nn.MPBaseReshape"):
        size = tuple(int(s) for s in o["size"].tolist())
        cls = {"nn.Reshape": tnn.Reshape, "nn.MPInitialReshape": tnn.MPInitialReshape,
               "nn.MPBaseReshape": tnn.MPBaseReshape}[c]
        return cls(*size, batchMode=o.get("batchMode"))
    if c == "nn.View":
        return tnn.View(*[int(s) for s in o["size"].tolist()])
    if c in ("nn.SpatialConvolutionMM", "nn.SpatialConvolution"):
        m = tnn.SpatialConvolutionMM(num("nInputPlane"), num("nOutputPlane"), num("kW"), num("kH"),
                                     num("dW", 1), num("dH", 1), num("padW", 0), num("padH", 0))
        with torch.no_grad():
            m.weight.copy_(o["weight"].float().reshape(m.weight.shape))
            m.bias.copy_(o["bias"].float())
        return m
    if c == "nn.SpatialMaxPooling":
        return tnn.SpatialMaxPooling(num("kW"), num("kH"), num("dW"), num("dH"), num("padW", 0), num("padH", 0))
    if c == "nn.Dropout":
        return tnn.Dropout(float(o.get("p", 0.5)))
    if c == "nn.LogSoftMax":
        return tnn.LogSoftMax()
    raise TypeError(f"no tadnn.nn counterpart for Torch7 class {c}")


def save_model(path: str, model: torch.nn.Module, mode: str = "binary") -> None:
    save(path, model_to_t7(model), mode)


def load_model(path: str, mode: str = "binary", mpi=None) -> torch.nn.Module:
    return model_from_t7(load(path, mode), mpi)
