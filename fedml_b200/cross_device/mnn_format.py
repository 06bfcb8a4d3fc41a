"""MNN model files (``.mnn``) without the MNN runtime: reader, in-place parameter writer and a graph builder.

A ``.mnn`` file is a FlatBuffers buffer whose root table is MNN's ``Net``; the reference's cross-device server only ever
(1) reads the trainable parameters of an uploaded file in graph order (cross_device/server_mnn/utils.py:11-28,
``module.parameters[i].read()``), (2) loads averaged values back into the same graph and saves it (:31-50), and
(3) creates the initial LeNet-5 / ResNet-20 graphs (model/mobile/mnn_lenet.py:31-35).  All three are done here on the
bytes directly:

* tables used and their field numbers (MNN schema): ``Net{oplists=3, tensorName=7}``, ``Op{inputIndexes=0, main_type=1,
  main=2, name=3, outputIndexes=4, type=5}``, ``Blob{dims=0, dataFormat=1, dataType=2, float32s=7}`` - a trainable
  parameter is an ``Op`` of type 266 (``TrainableParam``) whose ``main`` is a ``Blob`` (union tag 7);
* reading walks vtables; writing parameters overwrites the ``float32s`` payloads in place (sizes are checked), so every
  other byte of the graph - and therefore what the phones execute - stays as the device SDK wrote it;
* the builder emits a standard-conforming FlatBuffers buffer (vtable before table, children after parents, 4-byte
  alignment) for graphs made of Input / ConvertTensor / Convolution / ReLU / Pooling / Reshape / MatMul / BinaryOp /
  UnaryOp / Reduction / Const / Softmax / TrainableParam ops, which covers the reference's mobile models: LeNet-5, logistic
  regression and the ResNet-20 training graph (model/mobile/mnn_resnet.py:80-135), whose training-mode BatchNorm MNN
  serialises as mean / subtract / square / mean / +eps / sqrt / reciprocal / multiply / scale / shift primitives.
"""
from __future__ import annotations

import struct
from collections import OrderedDict
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

# OpType / OpParameter union tags of the ops handled here (MNN schema enums)
OP_BINARY, OP_CONV, OP_INPUT, OP_MATMUL, OP_POOLING, OP_RELU, OP_RESHAPE, OP_SOFTMAX = 7, 12, 34, 39, 47, 69, 73, 85
OP_CONVERT_TENSOR, OP_TRAINABLE_PARAM, OP_CONST, OP_REDUCTION, OP_UNARY = 129, 266, 11, 68, 101
PARAM_BINARY, PARAM_BLOB, PARAM_CONV2D, PARAM_INPUT, PARAM_MATMUL, PARAM_POOL, PARAM_RELU, PARAM_RESHAPE = 6, 7, 9, 21, 25, 31, 51, 55
PARAM_AXIS, PARAM_TENSOR_CONVERT, PARAM_REDUCTION, PARAM_UNARY = 4, 66, 50, 70
BINARY_ADD, BINARY_SUB, BINARY_MUL = 0, 1, 2
UNARY_SQUARE, UNARY_SQRT, UNARY_RSQRT, UNARY_RECIPROCAL = 4, 5, 6, 15
REDUCE_SUM, REDUCE_MEAN = 0, 3
FMT_NCHW, FMT_NHWC, FMT_NC4HW4 = 0, 1, 2
NET_OPLISTS, NET_TENSOR_NAME = 3, 7
OP_F_INPUTS, OP_F_MAIN_TYPE, OP_F_MAIN, OP_F_NAME, OP_F_OUTPUTS, OP_F_TYPE = 0, 1, 2, 3, 4, 5
BLOB_DIMS, BLOB_FORMAT, BLOB_DTYPE, BLOB_FLOATS = 0, 1, 2, 7


class MnnFormatError(ValueError):
    pass


# ------------------------------------------------------------------------------------------------ reading
class _Reader:
    def __init__(self, data: bytes):
        self.d = data
        if len(data) < 8:
            raise MnnFormatError("too short for a FlatBuffers buffer")

    def u32(self, o): return struct.unpack_from("<I", self.d, o)[0]
    def i32(self, o): return struct.unpack_from("<i", self.d, o)[0]
    def u16(self, o): return struct.unpack_from("<H", self.d, o)[0]

    def vtable(self, t) -> List[int]:
        vt = t - self.i32(t)
        if not 0 <= vt < len(self.d) - 4:
            raise MnnFormatError("vtable out of range")
        n = (self.u16(vt) - 4) // 2
        return [self.u16(vt + 4 + 2 * i) for i in range(n)]

    def field(self, t, i) -> Optional[int]:
        f = self.vtable(t)
        return t + f[i] if i < len(f) and f[i] else None

    def indirect(self, p): return p + self.u32(p)

    def vector(self, p) -> Tuple[int, int]:
        v = self.indirect(p)
        n = self.u32(v)
        if v + 4 > len(self.d):
            raise MnnFormatError("vector out of range")
        return v + 4, n

    def string(self, p) -> str:
        b, n = self.vector(p)
        return self.d[b:b + n].decode("utf-8", "replace")

    def ints(self, p) -> List[int]:
        b, n = self.vector(p)
        return list(struct.unpack_from(f"<{n}i", self.d, b))


def _ops(r: _Reader):
    root = r.u32(0)
    p = r.field(root, NET_OPLISTS)
    if p is None:
        raise MnnFormatError("Net.oplists missing: not an MNN model file")
    base, n = r.vector(p)
    for k in range(n):
        yield r.indirect(base + 4 * k)


def _param_slots(r: _Reader):
    """(name, dims, byte offset of the float payload, element count) of every TrainableParam op, in graph order."""
    out = []
    for op in _ops(r):
        tp = r.field(op, OP_F_TYPE)
        if tp is None or r.i32(tp) != OP_TRAINABLE_PARAM:
            continue
        mt, mp = r.field(op, OP_F_MAIN_TYPE), r.field(op, OP_F_MAIN)
        if mt is None or mp is None or r.d[mt] != PARAM_BLOB:
            raise MnnFormatError("TrainableParam without a Blob")
        blob = r.indirect(mp)
        dp, fp = r.field(blob, BLOB_DIMS), r.field(blob, BLOB_FLOATS)
        if fp is None:
            raise MnnFormatError("TrainableParam blob holds no float32 data")
        base, n = r.vector(fp)
        dims = r.ints(dp) if dp is not None else [n]
        if int(np.prod(dims)) != n or base + 4 * n > len(r.d):
            raise MnnFormatError(f"blob dims {dims} do not match {n} stored values")
        nm = r.field(op, OP_F_NAME)
        out.append((r.string(nm) if nm is not None else f"param{len(out)}", dims, base, n))
    return out


def read_params(path: str) -> "OrderedDict[str, np.ndarray]":
    """Trainable parameters of an MNN graph, graph order (= the ``module.parameters`` order of the reference's reader)."""
    with open(path, "rb") as fh:
        data = fh.read()
    r = _Reader(data)
    out = OrderedDict()
    for name, dims, base, n in _param_slots(r):
        out[name] = np.frombuffer(data, dtype="<f4", count=n, offset=base).reshape(dims).copy()
    return out


def write_params(path: str, values: Sequence[np.ndarray], out_path: Optional[str] = None) -> str:
    """Overwrites the float payload of the i-th trainable parameter with ``values[i]`` (element counts must match)."""
    with open(path, "rb") as fh:
        data = bytearray(fh.read())
    slots = _param_slots(_Reader(bytes(data)))
    if len(values) != len(slots):
        raise MnnFormatError(f"{path} has {len(slots)} trainable parameters, got {len(values)} tensors")
    for (name, dims, base, n), v in zip(slots, values):
        a = np.ascontiguousarray(np.asarray(v, dtype="<f4")).reshape(-1)
        if a.size != n:
            raise MnnFormatError(f"parameter {name}: file holds {n} values ({dims}), tensor has {a.size}")
        data[base:base + 4 * n] = a.tobytes()
    out_path = out_path or path
    tmp = out_path + ".tmp"
    with open(tmp, "wb") as fh:
        fh.write(data)
    import os
    os.replace(tmp, out_path)
    return out_path


def describe(path: str) -> List[dict]:
    """One dict per op: name, type, main_type, inputs, outputs (+ dims for parameters / inputs) - for tests and logging."""
    with open(path, "rb") as fh:
        r = _Reader(fh.read())
    ops = []
    for op in _ops(r):
        g = lambda i: r.field(op, i)                                  # noqa: E731
        e = {"name": r.string(g(OP_F_NAME)) if g(OP_F_NAME) else "", "type": r.i32(g(OP_F_TYPE)) if g(OP_F_TYPE) else 0,
             "main_type": r.d[g(OP_F_MAIN_TYPE)] if g(OP_F_MAIN_TYPE) else 0,
             "inputs": r.ints(g(OP_F_INPUTS)) if g(OP_F_INPUTS) else [], "outputs": r.ints(g(OP_F_OUTPUTS)) if g(OP_F_OUTPUTS) else []}
        if g(OP_F_MAIN):
            m = r.indirect(g(OP_F_MAIN))
            if e["main_type"] in (PARAM_BLOB, PARAM_INPUT, PARAM_RESHAPE) and r.field(m, 0) is not None:
                e["dims"] = r.ints(r.field(m, 0))
            if e["main_type"] == PARAM_CONV2D and r.field(m, 0) is not None:
                c = r.indirect(r.field(m, 0))
                iv = lambda i, dflt: r.i32(r.field(c, i)) if r.field(c, i) is not None else dflt      # noqa: E731
                e["conv"] = {"kernel": [iv(3, 1), iv(2, 1)], "stride": [iv(5, 1), iv(4, 1)], "out": iv(10, 0), "in": iv(11, 0), "pad": [iv(1, 0), iv(0, 0)],
                             "pad_mode": r.d[r.field(c, 8)] if r.field(c, 8) is not None else 0}
            if e["main_type"] == PARAM_POOL:
                iv = lambda i, dflt: r.i32(r.field(m, i)) if r.field(m, i) is not None else dflt      # noqa: E731
                e["pool"] = {"kernel": [iv(4, 0), iv(3, 0)], "stride": [iv(6, 0), iv(5, 0)]}
                e["pool_type"] = r.d[r.field(m, 7)] if r.field(m, 7) is not None else 0          # 0 max, 1 average
                e["pool_pad"] = r.d[r.field(m, 8)] if r.field(m, 8) is not None else 0           # 0 caffe, 1 valid, 2 same
            if e["main_type"] == PARAM_MATMUL:
                e["transpose_a"] = bool(r.d[r.field(m, 1)]) if r.field(m, 1) is not None else False
                e["transpose_b"] = bool(r.d[r.field(m, 2)]) if r.field(m, 2) is not None else False
            if e["main_type"] == PARAM_AXIS:
                e["axis"] = r.i32(r.field(m, 0)) if r.field(m, 0) is not None else 0
            if e["main_type"] == PARAM_BINARY:
                e["binary_op"] = r.i32(r.field(m, 0)) if r.field(m, 0) is not None else 0        # 0 add, 1 sub, 2 mul
            if e["main_type"] == PARAM_UNARY:
                e["unary_op"] = r.i32(r.field(m, 0)) if r.field(m, 0) is not None else 0         # 4 square, 5 sqrt, 15 1/x
            if e["main_type"] == PARAM_REDUCTION:
                e["reduce_op"] = r.d[r.field(m, 0)] if r.field(m, 0) is not None else 0          # 0 sum, 3 mean
                e["reduce_dims"] = r.ints(r.field(m, 1)) if r.field(m, 1) is not None else []
                e["keep_dims"] = bool(r.d[r.field(m, 3)]) if r.field(m, 3) is not None else False
            if e["main_type"] == PARAM_TENSOR_CONVERT:
                e["dest"] = r.d[r.field(m, 1)] if r.field(m, 1) is not None else 0
            if e["type"] == OP_CONST and e["main_type"] == PARAM_BLOB and r.field(m, BLOB_FLOATS) is not None:
                base, n = r.vector(r.field(m, BLOB_FLOATS))
                e["values"] = np.frombuffer(r.d, dtype="<f4", count=n, offset=base).copy()
        ops.append(e)
    return ops


# ------------------------------------------------------------------------------------------------ building
class _Builder:
    """Front-to-back FlatBuffers writer: a table is emitted with placeholder offsets, its children follow at higher
    addresses and the unsigned offsets are patched.  Field values: ("i32", v) | ("u8", v) | ("bool", v) | ("str", s) |
    ("ints", [..]) | ("floats", ndarray) | ("table", fields) | ("tables", [fields, ...]) | ("strs", [s, ...]);
    ``fields`` is a dict {field number: value}."""

    def __init__(self):
        self.b = bytearray(4)            # root offset, patched at the end

    def _align(self, n=4):
        while len(self.b) % n:
            self.b.append(0)

    def _child(self, kind, v) -> int:
        self._align()
        if kind == "str":
            raw = v.encode("utf-8")
            pos = len(self.b)
            self.b += struct.pack("<I", len(raw)) + raw + b"\0"
            return pos
        if kind == "ints":
            pos = len(self.b)
            self.b += struct.pack(f"<I{len(v)}i", len(v), *[int(x) for x in v])
            return pos
        if kind == "floats":
            a = np.ascontiguousarray(np.asarray(v, dtype="<f4")).reshape(-1)
            pos = len(self.b)
            self.b += struct.pack("<I", a.size) + a.tobytes()
            return pos
        if kind == "table":
            return self.table(v)
        if kind in ("tables", "strs"):
            pos = len(self.b)
            self.b += struct.pack("<I", len(v)) + bytes(4 * len(v))
            for i, item in enumerate(v):
                slot = pos + 4 + 4 * i
                cpos = self._child("table" if kind == "tables" else "str", item)
                struct.pack_into("<I", self.b, slot, cpos - slot)
            return pos
        raise ValueError(kind)

    def table(self, fields: Dict[int, tuple]) -> int:
        nf = (max(fields) + 1) if fields else 0
        vsize = 4 + 2 * nf
        self._align()
        if (len(self.b) + vsize) % 4:            # the table that follows the vtable must start 4-byte aligned
            self.b += b"\0\0"
        vpos = len(self.b)
        wide = [(i, k, v) for i, (k, v) in sorted(fields.items()) if k not in ("u8", "bool")]
        narrow = [(i, k, v) for i, (k, v) in sorted(fields.items()) if k in ("u8", "bool")]
        offs, cur = {}, 4
        for i, _, _ in wide:
            offs[i] = cur
            cur += 4
        for i, _, _ in narrow:
            offs[i] = cur
            cur += 1
        tsize = (cur + 3) // 4 * 4
        self.b += struct.pack("<HH", vsize, tsize) + b"".join(struct.pack("<H", offs.get(i, 0)) for i in range(nf))
        tpos = len(self.b)
        assert tpos % 4 == 0
        self.b += struct.pack("<i", tpos - vpos) + bytes(tsize - 4)
        pending = []
        for i, k, v in wide:
            if k == "i32":
                struct.pack_into("<i", self.b, tpos + offs[i], int(v))
            else:
                pending.append((tpos + offs[i], k, v))
        for i, k, v in narrow:
            self.b[tpos + offs[i]] = int(v) & 0xFF
        for slot, k, v in pending:
            cpos = self._child(k, v)
            struct.pack_into("<I", self.b, slot, cpos - slot)
        return tpos

    def finish(self, root_fields) -> bytes:
        root = self.table(root_fields)
        struct.pack_into("<I", self.b, 0, root)
        return bytes(self.b)


class MnnGraph:
    """Tiny graph DSL over the ops listed in the module docstring; every op produces one tensor named after the op
    (``<Kind><running number>``, the naming of MNN's own serialiser)."""

    def __init__(self):
        self.ops: List[dict] = []

    def _add(self, kind_name, op_type, main_type, main, inputs, name=None) -> int:
        idx = len(self.ops)
        self.ops.append({"name": name or f"{kind_name}{idx + 1}", "type": op_type, "main_type": main_type, "main": main,
                         "inputs": list(inputs), "outputs": [idx]})
        return idx

    def input(self, dims, fmt=FMT_NCHW):
        return self._add("Input", OP_INPUT, PARAM_INPUT, {0: ("ints", list(dims)), 2: ("u8", fmt)}, [])

    def convert(self, x, dest):
        return self._add("ConvertTensor", OP_CONVERT_TENSOR, PARAM_TENSOR_CONVERT, {1: ("u8", dest)} if dest else {}, [x])

    def param(self, array, name=None):
        a = np.asarray(array, dtype=np.float32)
        return self._add("TrainableParam", OP_TRAINABLE_PARAM, PARAM_BLOB, {BLOB_DIMS: ("ints", list(a.shape)), BLOB_FLOATS: ("floats", a)},
                         [], name)

    def conv(self, x, weight, bias, name=None, valid=True, stride=1, pad=0):
        """``valid`` False = MNN's SAME mode: output ceil(in / stride), padding split with the odd cell at the bottom / right."""
        w = np.asarray(weight)
        cout, cin, kh, kw = w.shape
        common = {2: ("i32", kw), 3: ("i32", kh), 8: ("u8", 1 if valid else 2), 10: ("i32", cout), 11: ("i32", cin)}
        if stride != 1:
            common.update({4: ("i32", stride), 5: ("i32", stride)})
        if pad:
            common.update({0: ("i32", pad), 1: ("i32", pad)})
        ins = [x, self.param(weight)]
        if bias is not None:
            ins.append(self.param(bias))
        return self._add("Conv", OP_CONV, PARAM_CONV2D, {0: ("table", common)}, ins, name)

    def const(self, value):
        a = np.asarray(value, dtype=np.float32)
        main = {1: ("u8", FMT_NHWC), 2: ("i32", 1), BLOB_FLOATS: ("floats", a)}
        if a.ndim:
            main[BLOB_DIMS] = ("ints", list(a.shape))
        return self._add("Const", OP_CONST, PARAM_BLOB, main, [])

    def binary(self, a, b, op=BINARY_ADD, name=None):
        return self._add("BinaryOp", OP_BINARY, PARAM_BINARY, {0: ("i32", op), 1: ("i32", 1)} if op else {1: ("i32", 1)}, [a, b], name)

    def unary(self, x, op):
        return self._add("UnaryOp", OP_UNARY, PARAM_UNARY, {0: ("i32", op), 1: ("i32", 1)}, [x])

    def reduce_mean(self, x, dims, keep_dims=True):
        return self._add("Reduction", OP_REDUCTION, PARAM_REDUCTION,
                         {0: ("u8", REDUCE_MEAN), 1: ("ints", list(dims)), 3: ("bool", int(keep_dims)), 4: ("i32", 1)}, [x])

    def avg_pool(self, x, k, s):
        return self._add("Pooling", OP_POOLING, PARAM_POOL,
                         {3: ("i32", k), 4: ("i32", k), 5: ("i32", s), 6: ("i32", s), 7: ("u8", 1), 8: ("u8", 1)}, [x])

    def batch_norm_train(self, x_nc4, scale, bias, name=None, eps=1e-5):
        """Training-mode BatchNorm the way MNN's ``nn.batch_norm`` serialises it: statistics of the batch over (N, H, W),
        biased variance, on an NCHW copy; the result goes back to NC4HW4."""
        x = self.convert(x_nc4, FMT_NCHW)
        centred = self.binary(x, self.reduce_mean(x, [0, 2, 3]), BINARY_SUB)
        var = self.reduce_mean(self.unary(centred, UNARY_SQUARE), [0, 2, 3])
        rstd = self.unary(self.unary(self.binary(var, self.const(eps)), UNARY_SQRT), UNARY_RECIPROCAL)
        y = self.binary(centred, rstd, BINARY_MUL)
        y = self.binary(y, self.param(np.asarray(scale, dtype=np.float32).reshape(1, -1, 1, 1)), BINARY_MUL)
        y = self.binary(y, self.param(np.asarray(bias, dtype=np.float32).reshape(1, -1, 1, 1)), BINARY_ADD, name)
        return self.convert(y, FMT_NC4HW4)

    def relu(self, x):
        return self._add("ReLU", OP_RELU, PARAM_RELU, {}, [x])

    def max_pool(self, x, k=2, s=2):
        return self._add("Pooling", OP_POOLING, PARAM_POOL, {3: ("i32", k), 4: ("i32", k), 5: ("i32", s), 6: ("i32", s), 8: ("u8", 1)}, [x])

    def reshape(self, x, dims):
        return self._add("Reshape", OP_RESHAPE, PARAM_RESHAPE, {0: ("ints", list(dims))}, [x])

    def linear(self, x, weight, bias):
        """y = x @ W^T + b with W [out, in] as MNN's nn.linear stores it (MatMul transposeB, bias [1, out])."""
        wi = self.param(weight)
        mm = self._add("MatMul", OP_MATMUL, PARAM_MATMUL, {2: ("bool", 1)}, [x, wi])
        bi = self.param(np.asarray(bias, dtype=np.float32).reshape(1, -1))
        return self.binary(mm, bi)

    def softmax(self, x, axis=1):
        return self._add("Softmax", OP_SOFTMAX, PARAM_AXIS, {0: ("i32", axis)}, [x])

    def to_bytes(self) -> bytes:
        fb = _Builder()
        ops = []
        for o in self.ops:
            f = {OP_F_MAIN_TYPE: ("u8", o["main_type"]), OP_F_MAIN: ("table", o["main"]), OP_F_NAME: ("str", o["name"]),
                 OP_F_OUTPUTS: ("ints", o["outputs"]), OP_F_TYPE: ("i32", o["type"])}
            if o["inputs"]:
                f[OP_F_INPUTS] = ("ints", o["inputs"])
            ops.append(f)
        return fb.finish({NET_OPLISTS: ("tables", ops), NET_TENSOR_NAME: ("strs", [o["name"] for o in self.ops])})

    def save(self, path: str) -> str:
        import os
        os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
        tmp = path + ".tmp"
        with open(tmp, "wb") as fh:
            fh.write(self.to_bytes())
        os.replace(tmp, path)
        return path


def build_lenet5(path: str, params: Sequence[np.ndarray], input_dims=(1, 1, 28, 28)) -> str:
    """The reference's mobile LeNet-5 (model/mobile/mnn_lenet.py:7-35): conv5x5(1->20) ReLU pool2 conv5x5(20->50) ReLU pool2,
    NC4HW4 -> NCHW, flatten, fc 800->500 ReLU, fc 500->classes, softmax.  ``params``: conv1.w, conv1.b, conv2.w, conv2.b,
    fc1.w, fc1.b, fc2.w, fc2.b (the order ``LeNet5().state_dict()`` gives)."""
    g = MnnGraph()
    x = g.convert(g.input(input_dims), FMT_NC4HW4)
    x = g.max_pool(g.relu(g.conv(x, params[0], params[1], name="conv1")))
    x = g.max_pool(g.relu(g.conv(x, params[2], params[3], name="conv2")))
    x = g.reshape(g.convert(x, FMT_NCHW), [0, -1])
    x = g.relu(g.linear(x, params[4], params[5]))
    g.softmax(g.linear(x, params[6], params[7]), 1)
    return g.save(path)


def build_mlp(path: str, params: Sequence[np.ndarray], input_dims) -> str:
    """flatten -> (linear, ReLU)* -> linear -> softmax; one (weight, bias) pair per layer.  One pair = the reference's
    logistic-regression mobile model."""
    g = MnnGraph()
    x = g.reshape(g.input(input_dims), [0, -1])
    n = len(params) // 2
    for i in range(n):
        x = g.linear(x, params[2 * i], params[2 * i + 1])
        if i + 1 < n:
            x = g.relu(x)
    g.softmax(x, 1)
    return g.save(path)


RESNET20_STAGES = ((16, 1), (16, 1), (16, 1), (32, 2), (32, 1), (32, 1), (64, 2), (64, 1), (64, 1))


def resnet20_param_shapes(class_num: int = 10, in_ch: int = 3) -> List[Tuple[int, ...]]:
    """Shapes of the 65 trainable tensors in graph order = ``named_parameters()`` order of the torch CIFAR ResNet-20."""
    shapes = [(16, in_ch, 3, 3), (16,), (16,)]
    cin = 16
    for cout, stride in RESNET20_STAGES:
        shapes += [(cout, cin, 3, 3), (cout,), (cout,), (cout, cout, 3, 3), (cout,), (cout,)]
        if stride != 1 or cin != cout:
            shapes += [(cout, cin, 1, 1), (cout,), (cout,)]
        cin = cout
    return shapes + [(class_num, 64), (class_num,)]


def build_resnet20(path: str, params: Sequence[np.ndarray], input_dims=(1, 3, 32, 32)) -> str:
    """The reference's mobile ResNet-20 (model/mobile/mnn_resnet.py:7-135) as MNN saves it from training mode: stem conv +
    BN + ReLU, nine two-convolution residual blocks (a 1x1 stride-2 convolution + BN on the shortcut where the width changes),
    8x8 average pool, flatten, fc, softmax.  Convolutions have no bias and use SAME padding; the additions run on NCHW
    tensors, so format conversions sit exactly where MNN's expression builder puts them.  ``params`` in
    :func:`resnet20_param_shapes` order (BN tensors may be flat)."""
    it = iter(params)
    g = MnnGraph()
    x = g.convert(g.input(input_dims), FMT_NC4HW4)
    x = g.relu(g.batch_norm_train(g.conv(x, next(it), None, "conv1", valid=False, pad=1), next(it), next(it), "bn1"))
    x_is_nc4, cin = True, 16
    for cout, stride in RESNET20_STAGES:
        xin = x if x_is_nc4 else g.convert(x, FMT_NC4HW4)
        y = g.conv(xin, next(it), None, "conv1", valid=False, stride=stride, pad=1)
        y = g.relu(g.batch_norm_train(y, next(it), next(it), "bn1"))
        y = g.batch_norm_train(g.conv(y, next(it), None, "conv2", valid=False, pad=1), next(it), next(it), "bn2")
        y = g.convert(y, FMT_NCHW)
        if stride != 1 or cin != cout:
            sc = g.conv(g.convert(x, FMT_NC4HW4), next(it), None, "conv_shortcut", valid=False, stride=stride)
            sc = g.convert(g.batch_norm_train(sc, next(it), next(it), "bn_shortcut"), FMT_NCHW)
        else:
            sc = g.convert(x, FMT_NCHW) if x_is_nc4 else x
        x = g.relu(g.binary(y, sc))
        x_is_nc4, cin = False, cout
    x = g.reshape(g.avg_pool(x, 8, 8), [0, -1])
    g.softmax(g.linear(x, next(it), next(it)), 1)
    return g.save(path)


def build_params_only(path: str, named: "OrderedDict[str, np.ndarray]") -> str:
    """A graph of TrainableParam ops named after the tensors - the container for architectures the op set above does not
    cover (the server only needs the parameters; a phone would need the full graph from its own SDK)."""
    g = MnnGraph()
    for k, v in named.items():
        g.param(np.asarray(v, dtype=np.float32), name=k)
    return g.save(path)


# ------------------------------------------------------------------------------------------------ executing
def run_graph(path: str, x, params: Optional[Sequence] = None):
    """Evaluates an MNN graph made of the ops listed in the module docstring on a torch tensor ``x`` (NCHW) and returns the
    output of the last op - the server-side evaluation the reference does with the MNN runtime
    (cross_device/server_mnn/fedml_aggregator.py:169-228), for the graphs this module can build and the LeNet-5 / logistic
    regression files a device SDK writes.  ``params``: optional replacement values for the trainable parameters, graph order.
    Raises :class:`MnnFormatError` on an op outside that set."""
    import torch
    import torch.nn.functional as F
    ops = describe(path)
    stored = iter(read_params(path).values()) if params is None else iter(params)
    vals: Dict[int, "torch.Tensor"] = {}
    out = None
    for o in ops:
        t, ins = o["type"], [vals[i] for i in o["inputs"]]
        if t == OP_INPUT:
            out = x.float()
        elif t == OP_TRAINABLE_PARAM:
            out = torch.as_tensor(np.asarray(next(stored)), dtype=torch.float32).reshape(o["dims"])
        elif t == OP_CONVERT_TENSOR:
            out = ins[0]                                   # NC4HW4 <-> NCHW is a storage matter of the device runtime
        elif t == OP_CONV:
            c = o["conv"]
            kh, kw = c["kernel"]
            xin = ins[0]
            if c["pad_mode"] == 2:                         # SAME: out = ceil(in / stride), the odd cell goes bottom / right
                sh, sw = c["stride"]
                ph = max((-(-xin.shape[2] // sh) - 1) * sh + kh - xin.shape[2], 0)
                pw = max((-(-xin.shape[3] // sw) - 1) * sw + kw - xin.shape[3], 0)
                xin = F.pad(xin, (pw // 2, pw - pw // 2, ph // 2, ph - ph // 2))
            elif c["pad_mode"] == 0 and (c.get("pad", [0, 0]) != [0, 0]):
                xin = F.pad(xin, (c["pad"][1], c["pad"][1], c["pad"][0], c["pad"][0]))
            out = F.conv2d(xin, ins[1], ins[2].reshape(-1) if len(ins) > 2 else None, stride=tuple(c["stride"]))
        elif t == OP_RELU:
            out = torch.relu(ins[0])
        elif t == OP_POOLING:
            pool = F.max_pool2d if o.get("pool_type", 0) == 0 else F.avg_pool2d
            out = pool(ins[0], tuple(o["pool"]["kernel"]), tuple(o["pool"]["stride"]))
        elif t == OP_RESHAPE:
            dims = [ins[0].shape[i] if d == 0 else d for i, d in enumerate(o["dims"])]
            out = ins[0].reshape(dims)
        elif t == OP_MATMUL:
            a, b = ins
            out = (a.transpose(-1, -2) if o.get("transpose_a") else a) @ (b.transpose(-1, -2) if o.get("transpose_b") else b)
        elif t == OP_BINARY:
            op = o.get("binary_op", 0)
            out = ins[0] + ins[1] if op == 0 else ins[0] - ins[1] if op == 1 else ins[0] * ins[1] if op == 2 else None
            if out is None:
                raise MnnFormatError(f"BinaryOp type {op} is not supported")
        elif t == OP_SOFTMAX:
            out = torch.softmax(ins[0], dim=o.get("axis", 1))
        elif t == OP_CONST:
            out = torch.as_tensor(o["values"], dtype=torch.float32).reshape(o.get("dims", []))
        elif t == OP_UNARY:
            fn = {UNARY_SQUARE: torch.square, UNARY_SQRT: torch.sqrt, UNARY_RSQRT: torch.rsqrt, UNARY_RECIPROCAL: torch.reciprocal,
                  0: torch.abs, 1: torch.neg, 7: torch.exp, 8: torch.log}.get(o["unary_op"])
            if fn is None:
                raise MnnFormatError(f"UnaryOp type {o['unary_op']} is not supported")
            out = fn(ins[0])
        elif t == OP_REDUCTION:
            fn = {REDUCE_SUM: torch.sum, REDUCE_MEAN: torch.mean}.get(o["reduce_op"])
            if fn is None:
                raise MnnFormatError(f"Reduction type {o['reduce_op']} is not supported")
            out = fn(ins[0], dim=tuple(o["reduce_dims"]) or None, keepdim=o["keep_dims"])
        else:
            raise MnnFormatError(f"op {o['name']!r} (type {t}) is outside the supported set")
        for idx in o["outputs"]:
            vals[idx] = out
    return out
