#!/usr/bin/env python
"""onnx_to_blob.py -- convert an NSNet2 baseline .onnx into the FVADNS2 weight blob libfvad_cuda reads.

    python tools/onnx_to_blob.py data/nsnet2-20ms-baseline.onnx data/nsnet2-20ms-baseline.fvadns2

The reference opens the .onnx with onnxruntime (src/NSNet2.zig:53-66); this library links no
onnxruntime and no `onnx` python package exists in the image, so this file carries its own
protobuf wire-format reader (varint / 64-bit / length-delimited / 32-bit) for the handful of ONNX
messages needed: ModelProto.graph, GraphProto.{node, initializer}, NodeProto, AttributeProto,
TensorProto.  Nothing about the network is assumed that the file can tell us:

  * the node list is walked from the graph input; the layer sequence must be
    FC [Relu] -> GRU -> GRU -> FC Relu -> FC Relu -> FC Sigmoid;
  * shape-only ops on the data path (Transpose, Squeeze, Unsqueeze, Reshape, Flatten, Identity, Cast) are
    EVALUATED on the concrete shape the reference feeds, [1, 54, 161] (src/NSNet2.zig:71-76): every FC
    must see [.., frames, features] and every GRU must see X = [seq = 54, batch = 1, in] (or
    [1, 54, in] with layout = 1), so a graph whose Transposes put the frame axis anywhere but the GRU
    sequence axis is refused instead of silently computing 54 one-step sequences;
  * GRU initial_h may be absent, an all-zero initializer, or the value a PyTorch export computes for
    h0 = None: a chain that ends in ConstantOfShape / Constant of value 0 (through Slice, Expand,
    Concat, ...) -- proven zero by walking its producers; anything else is refused (the reference binds
    no state tensor, NSNet2.zig:220, and the kernels start every chunk from h = 0);
  * FC weights may come from MatMul (stored [in][out], transposed here) or Gemm (transB honoured);
  * GRU: `linear_before_reset` must be 1, direction forward, activations default (Sigmoid, Tanh),
    W [1][3H][in], R [1][3H][H], B [1][6H] = Wb|Rb; ONNX gate order z, r, h is kept as is;
  * whether fc1 is followed by a Relu is recorded in the blob header (`fc1_relu`).

Anything else fails loudly.  `write_nsnet2_onnx()` writes a minimal .onnx of that topology from a
weight dict with the same wire-format code; tests use it to round-trip the converter.
"""
from __future__ import annotations

import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# ----------------------------------------------------------------------------- wire format: read


def _varint(buf, pos):
    out = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7
        if shift > 70:
            raise ValueError("protobuf: varint too long")


def fields(buf):
    """Yield (field_number, wire_type, value) of one message; value is int or memoryview."""
    buf = memoryview(buf)
    pos, n = 0, len(buf)
    while pos < n:
        key, pos = _varint(buf, pos)
        fno, wt = key >> 3, key & 7
        if wt == 0:
            v, pos = _varint(buf, pos)
        elif wt == 1:
            v = buf[pos:pos + 8]; pos += 8
        elif wt == 2:
            ln, pos = _varint(buf, pos)
            if pos + ln > n:
                raise ValueError("protobuf: truncated length-delimited field")
            v = buf[pos:pos + ln]; pos += ln
        elif wt == 5:
            v = buf[pos:pos + 4]; pos += 4
        else:
            raise ValueError(f"protobuf: unsupported wire type {wt}")
        yield fno, wt, v


def _packed_varints(v, wt):
    if wt == 0:
        return [v]
    out, pos = [], 0
    while pos < len(v):
        x, pos = _varint(v, pos)
        out.append(x)
    return out


def _s64(x):
    return x - (1 << 64) if x >= (1 << 63) else x


_DTYPES = {1: np.float32, 6: np.int32, 7: np.int64, 11: np.float64, 10: np.float16}


def parse_tensor(buf):
    dims, dtype, name, raw = [], 0, "", None
    fdata, idata, ddata = [], [], []
    for fno, wt, v in fields(buf):
        if fno == 1:
            dims += [_s64(x) for x in _packed_varints(v, wt)]
        elif fno == 2:
            dtype = v
        elif fno == 4:  # float_data (packed or not)
            fdata.append(np.frombuffer(v, "<f4"))
        elif fno == 5 or fno == 7:
            idata += [_s64(x) for x in _packed_varints(v, wt)]
        elif fno == 10:
            ddata.append(np.frombuffer(v, "<f8"))
        elif fno == 8:
            name = bytes(v).decode()
        elif fno == 9:
            raw = bytes(v)
        elif fno in (13, 14) and (fno == 13 or v == 1):
            raise ValueError(f"initializer {name!r}: external data is not supported")
    if dtype not in _DTYPES:
        raise ValueError(f"initializer {name!r}: unsupported data_type {dtype}")
    dt = np.dtype(_DTYPES[dtype]).newbyteorder("<")
    if raw is not None:
        arr = np.frombuffer(raw, dt)
    elif fdata:
        arr = np.concatenate(fdata)
    elif ddata:
        arr = np.concatenate(ddata)
    else:
        arr = np.asarray(idata, dtype=dt)
    cnt = int(np.prod(dims)) if dims else 1
    if arr.size != cnt:
        raise ValueError(f"initializer {name!r}: {arr.size} elements for dims {dims}")
    return name, arr.reshape(dims).astype(_DTYPES[dtype])


def parse_attribute(buf):
    name, val = "", None
    ints, floats, strings = [], [], []
    for fno, wt, v in fields(buf):
        if fno == 1:
            name = bytes(v).decode()
        elif fno == 2:
            val = struct.unpack("<f", v)[0]
        elif fno == 3:
            val = _s64(v)
        elif fno == 4:
            val = bytes(v)
        elif fno == 5:
            val = parse_tensor(v)[1]
        elif fno == 7:
            floats += list(np.frombuffer(v, "<f4")) if wt == 2 else [struct.unpack("<f", v)[0]]
        elif fno == 8:
            ints += [_s64(x) for x in _packed_varints(v, wt)]
        elif fno == 9:
            strings.append(bytes(v))
    if val is None:
        val = ints or floats or strings or None
    return name, val


def parse_node(buf):
    node = {"input": [], "output": [], "name": "", "op": "", "attr": {}}
    for fno, wt, v in fields(buf):
        if fno == 1:
            node["input"].append(bytes(v).decode())
        elif fno == 2:
            node["output"].append(bytes(v).decode())
        elif fno == 3:
            node["name"] = bytes(v).decode()
        elif fno == 4:
            node["op"] = bytes(v).decode()
        elif fno == 5:
            k, a = parse_attribute(v)
            node["attr"][k] = a
    return node


def parse_model(data: bytes):
    """-> (nodes, initializers{name: ndarray}, graph_inputs[names], graph_outputs[names])"""
    graph = None
    for fno, wt, v in fields(data):
        if fno == 7 and wt == 2:
            graph = v
    if graph is None:
        raise ValueError("not an ONNX ModelProto: no graph")
    nodes, inits, gin, gout = [], {}, [], []
    for fno, wt, v in fields(graph):
        if fno == 1:
            nodes.append(parse_node(v))
        elif fno == 5:
            k, a = parse_tensor(v)
            inits[k] = a
        elif fno in (11, 12):  # ValueInfoProto: name = 1
            for f2, _, v2 in fields(v):
                if f2 == 1:
                    (gin if fno == 11 else gout).append(bytes(v2).decode())
    # Constant nodes act as initializers
    for nd in nodes:
        if nd["op"] == "Constant" and "value" in nd["attr"] and isinstance(nd["attr"]["value"], np.ndarray):
            inits[nd["output"][0]] = nd["attr"]["value"]
    return nodes, inits, gin, gout


# ----------------------------------------------------------------------------- graph -> weights

# the tensor the reference binds as "input": [batch = 1, frames = 54, bins = 161] (src/NSNet2.zig:71-76)
REF_INPUT_SHAPE = (1, 54, 161)

# value-preserving ops that may sit between a zero constant and a GRU's initial_h
_ZERO_PASS = {"Expand", "Reshape", "Squeeze", "Unsqueeze", "Transpose", "Identity", "Cast", "Slice", "Tile", "Gather",
              "Flatten"}


def _is_zero(tensor):
    return tensor is None or not np.any(np.asarray(tensor))


def provably_zero(name, producers, inits, depth=0):
    """True when value `name` is zero whatever the graph input holds: an all-zero initializer / Constant,
    a ConstantOfShape whose fill value is 0 (its default), or such a value seen through ops that only
    move, select or replicate elements."""
    if depth > 64:
        return False
    if name in inits:
        return _is_zero(inits[name])
    nd = producers.get(name)
    if nd is None:
        return False
    op = nd["op"]
    if op == "ConstantOfShape":
        return _is_zero(nd["attr"].get("value"))
    if op == "Constant":
        v = nd["attr"].get("value")
        if v is None:
            v = nd["attr"].get("value_float", nd["attr"].get("value_floats", nd["attr"].get("value_int", nd["attr"].get("value_ints"))))
            if v is None:
                return False
        return _is_zero(v)
    if op in _ZERO_PASS:
        return bool(nd["input"]) and provably_zero(nd["input"][0], producers, inits, depth + 1)
    if op == "Concat":
        return bool(nd["input"]) and all(provably_zero(i, producers, inits, depth + 1) for i in nd["input"])
    if op == "Mul":
        return any(provably_zero(i, producers, inits, depth + 1) for i in nd["input"])
    return False


def _axes_of(nd, inits, what):
    """axes of Squeeze / Unsqueeze: attribute (opset < 13) or constant second input (opset >= 13)"""
    ax = nd["attr"].get("axes")
    if ax is None and len(nd["input"]) > 1 and nd["input"][1]:
        if nd["input"][1] not in inits:
            raise ValueError(f"{what} {nd['name']!r}: axes must be constant")
        ax = [int(x) for x in np.asarray(inits[nd["input"][1]]).reshape(-1)]
    return None if ax is None else [int(a) for a in (ax if isinstance(ax, (list, tuple)) else [ax])]


def shape_after(nd, shape, inits):
    """Concrete output shape of a shape-only op applied to a tensor of `shape` (no data is moved here: with
    batch = 1 the element order is decided by where the non-unit axes end up, which the callers check)."""
    op, rank = nd["op"], len(shape)
    if op in ("Identity", "Cast", "Dropout"):
        return shape
    if op == "Transpose":
        perm = nd["attr"].get("perm") or list(range(rank))[::-1]
        if sorted(perm) != list(range(rank)):
            raise ValueError(f"Transpose {nd['name']!r}: perm {perm} does not fit rank {rank}")
        return tuple(shape[p] for p in perm)
    if op == "Squeeze":
        ax = _axes_of(nd, inits, "Squeeze")
        ax = [i for i, d in enumerate(shape) if d == 1] if ax is None else [a % rank for a in ax]
        if any(shape[a] != 1 for a in ax):
            raise ValueError(f"Squeeze {nd['name']!r}: axis of size != 1 in {shape}")
        return tuple(d for i, d in enumerate(shape) if i not in ax)
    if op == "Unsqueeze":
        ax = _axes_of(nd, inits, "Unsqueeze")
        if ax is None:
            raise ValueError(f"Unsqueeze {nd['name']!r}: no axes")
        out = list(shape)
        for a in sorted(a % (rank + len(ax)) for a in ax):
            out.insert(a, 1)
        return tuple(out)
    if op == "Flatten":
        a = int(nd["attr"].get("axis", 1)) % (rank + 1)
        return (int(np.prod(shape[:a], dtype=np.int64)), int(np.prod(shape[a:], dtype=np.int64)))
    if op == "Reshape":
        if len(nd["input"]) < 2 or nd["input"][1] not in inits:
            raise ValueError(f"Reshape {nd['name']!r}: the target shape must be a constant")
        tgt = [int(x) for x in np.asarray(inits[nd["input"][1]]).reshape(-1)]
        out = [shape[i] if d == 0 else d for i, d in enumerate(tgt)]
        total = int(np.prod(shape, dtype=np.int64))
        if out.count(-1) > 1:
            raise ValueError(f"Reshape {nd['name']!r}: more than one -1")
        if -1 in out:
            known = int(np.prod([d for d in out if d != -1], dtype=np.int64))
            if known == 0 or total % known:
                raise ValueError(f"Reshape {nd['name']!r}: {shape} -> {tgt}")
            out[out.index(-1)] = total // known
        if int(np.prod(out, dtype=np.int64)) != total:
            raise ValueError(f"Reshape {nd['name']!r}: {shape} -> {tgt} changes the element count")
        return tuple(out)
    raise ValueError(f"unsupported op {op!r} ({nd['name']!r}) on the data path")


_SHAPE_OPS = {"Transpose", "Squeeze", "Unsqueeze", "Reshape", "Identity", "Flatten", "Cast", "Dropout"}


def _frames_by_features(shape, what):
    """The non-unit axes of an activation must be (frames, features) in this order: then its memory is the
    [54][features] matrix whatever unit axes surround them."""
    core = tuple(d for d in shape if d != 1)
    if len(core) != 2 or core[0] != REF_INPUT_SHAPE[1]:
        raise ValueError(f"{what}: activation shape {shape} is not [frames = {REF_INPUT_SHAPE[1]}][features] up to unit axes")
    return core[1]


def extract_weights(nodes, inits, gin, gout) -> dict:
    data_inputs = [n for n in gin if n not in inits]
    if len(data_inputs) != 1:
        raise ValueError(f"expected exactly one graph input, found {data_inputs}")
    consumers = {}
    for nd in nodes:
        for i in nd["input"]:
            consumers.setdefault(i, []).append(nd)

    producers = {o: nd for nd in nodes for o in nd["output"] if o}
    layers = []  # ("fc", w[out][in], b, act) / ("gru", W, R, Wb, Rb)
    cur = data_inputs[0]
    shape = REF_INPUT_SHAPE  # concrete shape of `cur` for the tensor the reference feeds
    guard = 0
    while cur not in gout:
        guard += 1
        if guard > 10_000:
            raise ValueError("graph walk did not terminate")
        nxt = [nd for nd in consumers.get(cur, []) if nd["op"] != "Shape"]
        if not nxt:
            raise ValueError(f"value {cur!r} has no consumer and is not a graph output")
        nd = nxt[0]
        op = nd["op"]
        if op == "MatMul":
            a, b = nd["input"]
            if b not in inits or a != cur:
                raise ValueError(f"MatMul {nd['name']!r}: expected activation x initializer")
            if _frames_by_features(shape, f"MatMul {nd['name']!r}") != inits[b].shape[0] or shape[-1] != inits[b].shape[0]:
                raise ValueError(f"MatMul {nd['name']!r}: activation {shape} x weight {inits[b].shape}")
            layers.append(["fc", inits[b].T.copy(), None, None])
            shape = shape[:-1] + (int(inits[b].shape[1]),)
        elif op == "Gemm":
            a, b = nd["input"][:2]
            if b not in inits or a != cur or nd["attr"].get("transA", 0):
                raise ValueError(f"Gemm {nd['name']!r}: expected activation x initializer")
            w = inits[b] * np.float32(nd["attr"].get("alpha", 1.0))
            w = w.copy() if nd["attr"].get("transB", 0) else w.T.copy()
            bias = None
            if len(nd["input"]) > 2:
                bias = inits[nd["input"][2]].reshape(-1) * np.float32(nd["attr"].get("beta", 1.0))
            if len(shape) != 2 or _frames_by_features(shape, f"Gemm {nd['name']!r}") != w.shape[1]:
                raise ValueError(f"Gemm {nd['name']!r}: activation {shape} x weight [out][in] {w.shape}")
            layers.append(["fc", w, bias, None])
            shape = (shape[0], int(w.shape[0]))
        elif op == "Add":
            other = [i for i in nd["input"] if i != cur]
            if len(other) != 1 or other[0] not in inits or not layers or layers[-1][0] != "fc" or layers[-1][2] is not None:
                raise ValueError(f"Add {nd['name']!r}: only a bias after MatMul is understood")
            layers[-1][2] = inits[other[0]].reshape(-1).copy()
        elif op in ("Relu", "Sigmoid"):
            if not layers or layers[-1][0] != "fc" or layers[-1][3] is not None:
                raise ValueError(f"{op} {nd['name']!r}: activation without a preceding FC")
            layers[-1][3] = op
        elif op == "GRU":
            at = nd["attr"]
            if at.get("linear_before_reset", 0) != 1:
                raise ValueError("GRU: linear_before_reset must be 1 (kernels implement r * (R_h h + Rb_h))")
            if at.get("direction", b"forward") not in (b"forward", "forward"):
                raise ValueError("GRU: only direction=forward")
            if at.get("activations") not in (None, [b"Sigmoid", b"Tanh"]):
                raise ValueError(f"GRU: unsupported activations {at.get('activations')}")
            ins = nd["input"] + [""] * 6
            if ins[0] != cur:
                raise ValueError("GRU: X is not the running activation")
            if ins[4]:
                raise ValueError("GRU: sequence_lens is not supported")
            if ins[5] and not provably_zero(ins[5], producers, inits):
                raise ValueError("GRU: initial_h must be absent or provably zero (the reference binds no state tensor, "
                                 "NSNet2.zig:220; the kernels start every chunk from h = 0)")
            if at.get("clip") is not None:
                raise ValueError("GRU: clip is not supported")
            W, R = inits[ins[1]], inits[ins[2]]
            if W.ndim != 3 or W.shape[0] != 1 or R.shape[0] != 1:
                raise ValueError("GRU: expected W [1][3H][in], R [1][3H][H]")
            H = int(at.get("hidden_size", R.shape[2]))
            if W.shape[1] != 3 * H or R.shape[1:] != (3 * H, H):
                raise ValueError("GRU: W/R shapes disagree with hidden_size")
            if ins[3]:
                B = inits[ins[3]].reshape(-1)
                if B.size != 6 * H:
                    raise ValueError("GRU: B must be [1][6H]")
                Wb, Rb = B[:3 * H].copy(), B[3 * H:].copy()
            else:
                Wb, Rb = np.zeros(3 * H, np.float32), np.zeros(3 * H, np.float32)
            # X must be [seq = frames, batch = 1, in] (layout 0) or [1, frames, in] (layout 1): the frame axis IS the
            # recurrence axis.  A graph that feeds [1, 54, in] with layout 0 would run 54 one-step sequences.
            T, lay = REF_INPUT_SHAPE[1], int(at.get("layout", 0))
            want = (T, 1, int(W.shape[2])) if lay == 0 else (1, T, int(W.shape[2]))
            if tuple(shape) != want:
                raise ValueError(f"GRU {nd['name']!r}: X has shape {tuple(shape)}, expected {want} "
                                 f"(layout = {lay}): the frame axis must be the GRU sequence axis")
            layers.append(["gru", W[0].copy(), R[0].copy(), Wb, Rb])
            shape = (T, 1, 1, H) if lay == 0 else (1, T, 1, H)  # Y: [seq, num_directions, batch, H] / [batch, seq, dirs, H]
        elif op in _SHAPE_OPS:
            shape = shape_after(nd, shape, inits)
            _frames_by_features(shape, f"{op} {nd['name']!r}")
        else:
            raise ValueError(f"unsupported op {op!r} ({nd['name']!r}) on the data path")
        # GRU outputs: Y (all steps) is the one that continues
        outs = [o for o in nd["output"] if o and (o in consumers or o in gout)]
        if not outs or (op == "GRU" and outs[0] != nd["output"][0]):
            raise ValueError(f"node {nd['name']!r}: no consumed output (a GRU must continue through Y)")
        cur = outs[0]
    if tuple(shape) != (REF_INPUT_SHAPE[0], REF_INPUT_SHAPE[1], REF_INPUT_SHAPE[2]):
        raise ValueError(f"graph output has shape {tuple(shape)}, the reference reads {REF_INPUT_SHAPE} (NSNet2.zig:99-106)")

    kinds = [l[0] for l in layers]
    if kinds != ["fc", "gru", "gru", "fc", "fc", "fc"]:
        raise ValueError(f"layer sequence {kinds} is not FC, GRU, GRU, FC, FC, FC")
    fc1, g1, g2, fc2, fc3, fc4 = layers
    if fc1[3] not in (None, "Relu") or fc2[3] != "Relu" or fc3[3] != "Relu" or fc4[3] != "Sigmoid":
        raise ValueError(f"activations {[fc1[3], fc2[3], fc3[3], fc4[3]]} are not [None|Relu, Relu, Relu, Sigmoid]")
    for l in (fc1, fc2, fc3, fc4):
        if l[2] is None:
            l[2] = np.zeros(l[1].shape[0], np.float32)
    f32 = lambda a: np.ascontiguousarray(a, np.float32)
    wd = {"fc1_relu": int(fc1[3] == "Relu"),
          "fc1_w": f32(fc1[1]), "fc1_b": f32(fc1[2]),
          "gru1_W": f32(g1[1]), "gru1_R": f32(g1[2]), "gru1_Wb": f32(g1[3]), "gru1_Rb": f32(g1[4]),
          "gru2_W": f32(g2[1]), "gru2_R": f32(g2[2]), "gru2_Wb": f32(g2[3]), "gru2_Rb": f32(g2[4]),
          "fc2_w": f32(fc2[1]), "fc2_b": f32(fc2[2]), "fc3_w": f32(fc3[1]), "fc3_b": f32(fc3[2]),
          "fc4_w": f32(fc4[1]), "fc4_b": f32(fc4[2])}
    H, B = wd["fc1_w"].shape
    import importlib
    shapes = importlib.import_module("fvad_b200").load().weights.shapes(B, H, wd["fc2_w"].shape[0], wd["fc3_w"].shape[0])
    for k, shp in shapes.items():
        if wd[k].shape != shp:
            raise ValueError(f"{k}: shape {wd[k].shape}, expected {shp}")
    return wd


def convert(onnx_path: str, blob_path: str | None = None) -> dict:
    with open(onnx_path, "rb") as f:
        data = f.read()
    wd = extract_weights(*parse_model(data))
    if blob_path:
        import fvad_b200
        fvad_b200.load().weights.save_blob(blob_path, wd)
    return wd


# ----------------------------------------------------------------------------- wire format: write


def _enc_varint(x: int) -> bytes:
    x &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = x & 0x7F
        x >>= 7
        out.append(b | (0x80 if x else 0))
        if not x:
            return bytes(out)


def _ld(fno: int, payload: bytes) -> bytes:
    return _enc_varint(fno << 3 | 2) + _enc_varint(len(payload)) + payload


def _vi(fno: int, x: int) -> bytes:
    return _enc_varint(fno << 3 | 0) + _enc_varint(x)


def _tensor(name: str, arr: np.ndarray, raw: bool = True) -> bytes:
    arr = np.ascontiguousarray(arr, np.float32)
    out = b"".join(_vi(1, d) for d in arr.shape) + _vi(2, 1)
    out += _ld(9, arr.tobytes()) if raw else _ld(4, arr.tobytes())
    return out + _ld(8, name.encode())


def _tensor_i64(name: str, vals) -> bytes:
    arr = np.ascontiguousarray(vals, np.int64)
    return b"".join(_vi(1, d) for d in arr.shape) + _vi(2, 7) + _ld(9, arr.tobytes()) + _ld(8, name.encode())


def _attr_i(name, v):
    return _ld(1, name.encode()) + _vi(3, v) + _vi(20, 2)


def _attr_ints(name, vs):
    return _ld(1, name.encode()) + b"".join(_vi(8, v) for v in vs) + _vi(20, 7)


def _attr_t(name, tensor_bytes):
    return _ld(1, name.encode()) + _ld(5, tensor_bytes) + _vi(20, 4)


def _node(op, ins, outs, name="", attrs=()):
    out = b"".join(_ld(1, i.encode()) for i in ins) + b"".join(_ld(2, o.encode()) for o in outs)
    return out + _ld(3, name.encode()) + _ld(4, op.encode()) + b"".join(_ld(5, a) for a in attrs)


def write_nsnet2_onnx(path: str, wd: dict, gemm: bool = False, raw: bool = True, torch_style: bool = False,
                      drop_transposes: bool = False, h0_value: float = 0.0) -> None:
    """Minimal NSNet2-baseline-shaped .onnx (test fixture writer).

    Flavours: MatMul+Add on the rank-3 tensor (default), Gemm on the squeezed [54][features] matrix
    (`gemm`), and `torch_style` -- what torch.onnx.export emits for a batch_first model whose GRUs get
    h0 = None: Transpose(1,0,2) around the recurrent block, initial_h = Slice(ConstantOfShape(Concat(
    [layers], Gather(Shape(x), 1), [H]))) per layer, Squeeze with `axes` as an input (opset 13).
    `drop_transposes` and `h0_value` produce graphs the importer must refuse."""
    nodes, inits = [], []

    def fc(x, name, w, b, act):
        if gemm:
            inits.append(_tensor(name + ".weight", w, raw)); inits.append(_tensor(name + ".bias", b, raw))
            nodes.append(_node("Gemm", [x, name + ".weight", name + ".bias"], [name + "_o"], name, [_attr_i("transB", 1)]))
        else:
            inits.append(_tensor(name + ".wT", np.asarray(w).T, raw)); inits.append(_tensor(name + ".bias", b, raw))
            nodes.append(_node("MatMul", [x, name + ".wT"], [name + "_mm"], name + "_mm"))
            nodes.append(_node("Add", [name + ".bias", name + "_mm"], [name + "_o"], name + "_add"))
        y = name + "_o"
        if act:
            nodes.append(_node(act, [y], [name + "_a"], name + "_act"))
            y = name + "_a"
        return y

    def squeeze(x, out, axes, name):
        if torch_style:  # opset 13: axes is an input
            inits.append(_tensor_i64(name + ".axes", axes))
            nodes.append(_node("Squeeze", [x, name + ".axes"], [out], name))
        else:
            nodes.append(_node("Squeeze", [x], [out], name, [_attr_ints("axes", axes)]))
        return out

    def gru(x, name, W, R, Wb, Rb, h0=None):
        H = R.shape[1]
        inits.append(_tensor(name + ".W", W[None], raw)); inits.append(_tensor(name + ".R", R[None], raw))
        inits.append(_tensor(name + ".B", np.concatenate([Wb, Rb])[None], raw))
        ins = [x, name + ".W", name + ".R", name + ".B"] + (["", h0] if h0 else [])
        nodes.append(_node("GRU", ins, [name + "_Y", name + "_Yh"], name,
                           [_attr_i("hidden_size", H), _attr_i("linear_before_reset", 1)]))
        return squeeze(name + "_Y", name + "_sq", [1], name + "_squeeze")

    x = "input"
    if gemm:  # Gemm is rank 2: [1][54][161] -> [54][161]
        x = squeeze(x, "in2d", [0], "in_squeeze")
    y = fc(x, "fc1", wd["fc1_w"], wd["fc1_b"], "Relu" if wd.get("fc1_relu") else None)
    if gemm:
        nodes.append(_node("Unsqueeze", [y], ["tr0"], "unsq0", [_attr_ints("axes", [1])]))  # [54][1][400]
    elif drop_transposes:
        nodes.append(_node("Identity", [y], ["tr0"], "tr0"))
    else:
        nodes.append(_node("Transpose", [y], ["tr0"], "tr0", [_attr_ints("perm", [1, 0, 2])]))
    h0 = [None, None]
    if torch_style:
        # h0 = zeros(num_layers, batch, H), batch read from the running tensor's shape, then one slice per layer
        H = wd["gru1_R"].shape[1]
        nodes.append(_node("Shape", ["tr0"], ["h0_shape"], "h0_shape"))
        inits.append(_tensor_i64("h0_one", [1]))
        nodes.append(_node("Gather", ["h0_shape", "h0_one"], ["h0_batch"], "h0_gather", [_attr_i("axis", 0)]))
        inits.append(_tensor_i64("h0_layers", [2])); inits.append(_tensor_i64("h0_hidden", [H]))
        nodes.append(_node("Concat", ["h0_layers", "h0_batch", "h0_hidden"], ["h0_dims"], "h0_concat", [_attr_i("axis", 0)]))
        fill = _vi(1, 1) + _vi(2, 1) + _ld(9, np.float32(h0_value).tobytes())
        nodes.append(_node("ConstantOfShape", ["h0_dims"], ["h0"], "h0_fill", [_attr_t("value", fill)]))
        for l in range(2):
            inits.append(_tensor_i64(f"h0_s{l}", [l])); inits.append(_tensor_i64(f"h0_e{l}", [l + 1]))
            inits.append(_tensor_i64(f"h0_ax{l}", [0]))
            nodes.append(_node("Slice", ["h0", f"h0_s{l}", f"h0_e{l}", f"h0_ax{l}"], [f"h0_l{l}"], f"h0_slice{l}"))
            h0[l] = f"h0_l{l}"
    y = gru("tr0", "rnn1", wd["gru1_W"], wd["gru1_R"], wd["gru1_Wb"], wd["gru1_Rb"], h0[0])
    y = gru(y, "rnn2", wd["gru2_W"], wd["gru2_R"], wd["gru2_Wb"], wd["gru2_Rb"], h0[1])
    if gemm:
        y = squeeze(y, "tr1", [1], "sq1")  # [54][400]
    elif drop_transposes:
        nodes.append(_node("Identity", [y], ["tr1"], "tr1")); y = "tr1"
    else:
        nodes.append(_node("Transpose", [y], ["tr1"], "tr1", [_attr_ints("perm", [1, 0, 2])])); y = "tr1"
    y = fc(y, "fc2", wd["fc2_w"], wd["fc2_b"], "Relu")
    y = fc(y, "fc3", wd["fc3_w"], wd["fc3_b"], "Relu")
    y = fc(y, "fc4", wd["fc4_w"], wd["fc4_b"], "Sigmoid")
    if gemm:
        nodes.append(_node("Unsqueeze", [y], ["output"], "out", [_attr_ints("axes", [0])]))
    else:
        nodes.append(_node("Identity", [y], ["output"], "out"))
    graph = b"".join(_ld(1, n) for n in nodes) + _ld(2, b"nsnet2") + b"".join(_ld(5, t) for t in inits)
    graph += _ld(11, _ld(1, b"input")) + _ld(12, _ld(1, b"output"))
    model = _vi(1, 7) + _ld(2, b"fvad-b200-test") + _ld(7, graph)
    with open(path, "wb") as f:
        f.write(model)


if __name__ == "__main__":
    if len(sys.argv) != 3:
        raise SystemExit(__doc__)
    w = convert(sys.argv[1], sys.argv[2])
    print(f"wrote {sys.argv[2]}: fc1_relu={w['fc1_relu']} hidden={w['fc1_w'].shape[0]} bins={w['fc1_w'].shape[1]}")
