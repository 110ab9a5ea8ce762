"""Minimal protobuf wire codec for the ONNX messages the bridge needs.

The ``onnx`` python package is not part of this image, so the bridge serialises ModelProto /
GraphProto / NodeProto / TensorProto / ValueInfoProto / AttributeProto itself (field numbers from
onnx.proto3, IR version 8).  Files written here load in stock onnx / onnxruntime and vice versa.
"""
from __future__ import annotations

import struct
from typing import Dict, List, Tuple

import numpy as np

FLOAT, INT32, INT64, BOOL = 1, 6, 7, 9
_NP = {FLOAT: np.float32, INT32: np.int32, INT64: np.int64, 11: np.float64, 9: np.bool_, 10: np.float16}
A_FLOAT, A_INT, A_STRING, A_TENSOR, A_FLOATS, A_INTS = 1, 2, 3, 4, 6, 7


# ------------------------------------------------------------------------------ writer
def _varint(v: int) -> bytes:
    if v < 0:
        v += 1 << 64
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _key(field: int, wt: int) -> bytes:
    return _varint((field << 3) | wt)


def w_int(field, v):
    return _key(field, 0) + _varint(int(v))


def w_bytes(field, b):
    if isinstance(b, str):
        b = b.encode()
    return _key(field, 2) + _varint(len(b)) + b


def w_float(field, v):
    return _key(field, 5) + struct.pack('<f', float(v))


def w_packed_ints(field, vals):
    body = b''.join(_varint(int(v)) for v in vals)
    return _key(field, 2) + _varint(len(body)) + body


def tensor(name: str, arr: np.ndarray) -> bytes:
    arr = np.ascontiguousarray(arr)
    dt = {np.dtype(np.float32): FLOAT, np.dtype(np.int64): INT64, np.dtype(np.int32): INT32}.get(arr.dtype)
    if dt is None:
        arr, dt = arr.astype(np.float32), FLOAT
    return (w_packed_ints(1, arr.shape) if arr.ndim else b'') + w_int(2, dt) + w_bytes(8, name) + w_bytes(9, arr.tobytes())


def attribute(name: str, value) -> bytes:
    out = w_bytes(1, name)
    if isinstance(value, bool):
        out += w_int(3, int(value)) + w_int(20, A_INT)
    elif isinstance(value, int):
        out += w_int(3, value) + w_int(20, A_INT)
    elif isinstance(value, float):
        out += w_float(2, value) + w_int(20, A_FLOAT)
    elif isinstance(value, str):
        out += w_bytes(4, value) + w_int(20, A_STRING)
    elif isinstance(value, np.ndarray):
        out += w_bytes(5, tensor('', value)) + w_int(20, A_TENSOR)
    elif isinstance(value, (list, tuple)) and all(isinstance(v, (int, np.integer)) for v in value):
        out += w_packed_ints(8, value) + w_int(20, A_INTS)
    elif isinstance(value, (list, tuple)):
        body = b''.join(struct.pack('<f', float(v)) for v in value)
        out += _key(7, 2) + _varint(len(body)) + body + w_int(20, A_FLOATS)
    else:
        raise TypeError('attribute %s: %r' % (name, type(value)))
    return out


def node(op_type: str, inputs: List[str], outputs: List[str], name: str = '', **attrs) -> bytes:
    out = b''.join(w_bytes(1, i) for i in inputs) + b''.join(w_bytes(2, o) for o in outputs)
    out += w_bytes(3, name or outputs[0]) + w_bytes(4, op_type)
    for k, v in attrs.items():
        out += w_bytes(5, attribute(k, v))
    return out


def value_info(name: str, shape, elem_type=FLOAT) -> bytes:
    dims = b''
    for d in shape:
        dims += w_bytes(1, w_int(1, d) if isinstance(d, (int, np.integer)) and d >= 0 else w_bytes(2, str(d)))
    ttype = w_int(1, elem_type) + w_bytes(2, dims)
    return w_bytes(1, name) + w_bytes(2, w_bytes(1, ttype))


def model(nodes: List[bytes], inputs: List[bytes], outputs: List[bytes], initializers: List[bytes],
          graph_name='hetu_graph', opset=17) -> bytes:
    g = b''.join(w_bytes(1, n) for n in nodes) + w_bytes(2, graph_name)
    g += b''.join(w_bytes(5, t) for t in initializers)
    g += b''.join(w_bytes(11, i) for i in inputs) + b''.join(w_bytes(12, o) for o in outputs)
    return (w_int(1, 8) + w_bytes(2, 'hetu_61a7_b200') + w_bytes(3, '0.1') + w_bytes(7, g)
            + w_bytes(8, w_bytes(1, '') + w_int(2, opset)))


# ------------------------------------------------------------------------------ reader
def parse(buf: bytes) -> Dict[int, list]:
    """One message level: {field: [(wire_type, value), ...]} (length-delimited values stay bytes)."""
    out: Dict[int, list] = {}
    i, n = 0, len(buf)
    while i < n:
        key, i = _read_varint(buf, i)
        field, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _read_varint(buf, i)
        elif wt == 1:
            v, i = buf[i:i + 8], i + 8
        elif wt == 2:
            ln, i = _read_varint(buf, i)
            v, i = buf[i:i + ln], i + ln
        elif wt == 5:
            v, i = buf[i:i + 4], i + 4
        else:
            raise ValueError('unsupported wire type %d' % wt)
        out.setdefault(field, []).append((wt, v))
    return out


def _read_varint(buf, i) -> Tuple[int, int]:
    shift = v = 0
    while True:
        b = buf[i]
        i += 1
        v |= (b & 0x7F) << shift
        if not b & 0x80:
            break
        shift += 7
    return v, i


def _signed(v):
    return v - (1 << 64) if v >= (1 << 63) else v


def _ints(entries) -> List[int]:
    out = []
    for wt, v in entries:
        if wt == 0:
            out.append(_signed(v))
        else:                                   # packed
            i = 0
            while i < len(v):
                x, i = _read_varint(v, i)
                out.append(_signed(x))
    return out


def read_tensor(buf: bytes):
    m = parse(buf)
    dims = _ints(m.get(1, []))
    dt = m[2][0][1] if 2 in m else FLOAT
    name = m[8][0][1].decode() if 8 in m else ''
    npdt = _NP.get(dt, np.float32)
    if 9 in m:
        arr = np.frombuffer(m[9][0][1], dtype=npdt).copy()
    elif 4 in m:                                # float_data
        raw = b''.join(v for _, v in m[4])
        arr = np.frombuffer(raw, dtype=np.float32).copy()
    elif 7 in m:
        arr = np.array(_ints(m[7]), dtype=np.int64)
    elif 5 in m:
        arr = np.array(_ints(m[5]), dtype=np.int32)
    else:
        arr = np.zeros(0, dtype=npdt)
    return name, arr.reshape(dims) if dims else arr.reshape(())


def read_attribute(buf: bytes):
    m = parse(buf)
    name = m[1][0][1].decode()
    typ = m[20][0][1] if 20 in m else 0
    if typ == A_FLOAT or (typ == 0 and 2 in m):
        return name, struct.unpack('<f', m[2][0][1])[0]
    if typ == A_INT or (typ == 0 and 3 in m):
        return name, _signed(m[3][0][1])
    if typ == A_STRING or (typ == 0 and 4 in m):
        return name, m[4][0][1].decode()
    if typ == A_TENSOR or (typ == 0 and 5 in m):
        return name, read_tensor(m[5][0][1])[1]
    if typ == A_INTS or (typ == 0 and 8 in m):
        return name, _ints(m.get(8, []))
    if typ == A_FLOATS or (typ == 0 and 7 in m):
        raw = b''.join(v for _, v in m.get(7, []))
        return name, list(np.frombuffer(raw, dtype=np.float32))
    return name, None


def read_value_info(buf: bytes):
    m = parse(buf)
    name = m[1][0][1].decode()
    shape, elem = [], FLOAT
    if 2 in m:
        t = parse(m[2][0][1])
        if 1 in t:
            tt = parse(t[1][0][1])
            elem = tt[1][0][1] if 1 in tt else FLOAT
            if 2 in tt:
                for _, d in parse(tt[2][0][1]).get(1, []):
                    dm = parse(d)
                    shape.append(_signed(dm[1][0][1]) if 1 in dm else -1)
    return name, shape, elem


def read_model(buf: bytes):
    """-> dict(nodes=[(op_type, inputs, outputs, attrs, name)], inputs, outputs, initializers, opset)."""
    m = parse(buf)
    g = parse(m[7][0][1])
    nodes = []
    for _, nb in g.get(1, []):
        nm = parse(nb)
        attrs = dict(read_attribute(a) for _, a in nm.get(5, []))
        nodes.append((nm[4][0][1].decode(), [v.decode() for _, v in nm.get(1, [])],
                      [v.decode() for _, v in nm.get(2, [])], attrs, nm[3][0][1].decode() if 3 in nm else ''))
    inits = dict(read_tensor(t) for _, t in g.get(5, []))
    opset = 0
    for _, o in m.get(8, []):
        om = parse(o)
        opset = max(opset, om[2][0][1] if 2 in om else 0)
    return dict(nodes=nodes, inputs=[read_value_info(v) for _, v in g.get(11, [])],
                outputs=[read_value_info(v) for _, v in g.get(12, [])], initializers=inits, opset=opset)
