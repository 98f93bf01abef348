#!/usr/bin/env python3
"""Convert an NNoM-generated ``weights.h`` whose graph is NOT the plain conv/dense chain of the reference model into a
version-2 ("graph") blob for libedison_b200's generic executor and the oracle.

Vocabulary (what nnom_utils.py:447-745 emits and this NNoM version runs reproducibly): Input, Conv2D (any kernel, stride,
PADDING_VALID / PADDING_SAME), act_relu tails (model.active), MaxPool (any kernel, stride, padding), GlobalMaxPool,
Concat(-1) of equally shaped inputs (model.mergex), Add (model.mergex, with its output shift), Dense, Softmax, Output.
The header's layer[] indices become node indices; a model.active() line aliases its source layer (NNoM attaches the
activation to that layer's tail, nnom.c model_active -- whoever hooks to either index reads the activated output).

Blob layout (little endian):
  char[8] "EDNNQ7\\0\\0", u32 version = 2, u32 n_nodes
  per node: u32 op, u32 n_in, u32 in[8], u32 p[10], then for conv/dense: int8 w[], int8 b[], zero padding to 4 bytes
     op 0 input    p = h, w, c
     op 1 conv     p = cout, kh, kw, stride_h, stride_w, same, out_rshift, bias_lshift, relu, cin     w = [cout][kh][kw][cin]
     op 2 dense    p = cout, cin, out_rshift, bias_lshift, relu                                       w = [cout][cin] row major
     op 3 maxpool  p = kh, kw, stride_h, stride_w, same
     op 4 concat   (channel axis)
     op 5 softmax
     op 6 global max pool
     op 7 add      p = out_rshift, relu
  trailer: i32 input_scale, i32 input_min, i32 input_max
usage: convert_nnom_graph.py <weights.h> <out.bin>
"""
import re
import struct
import sys

import numpy as np

sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.abspath(__file__)))
from convert_nnom_weights import deinterleave_x4  # noqa: E402

OP = dict(input=0, conv=1, dense=2, maxpool=3, concat=4, softmax=5, gmaxpool=6, add=7)
NONE = 0xFFFFFFFF


class Header:
    def __init__(self, path):
        self.txt = open(path).read()
        self.defs = {m.group(1): m.group(2).strip() for m in re.finditer(r"^#define\s+(\w+)\s+(.+?)\s*$", self.txt, re.M)}

    def value(self, expr):
        """integer value of a macro expression (names, + - * parentheses)"""
        expr = expr.strip()
        for _ in range(64):
            names = [n for n in re.findall(r"[A-Za-z_]\w*", expr)]
            if not names:
                break
            for n in names:
                expr = re.sub(r"\b%s\b" % n, "(" + self.defs[n] + ")", expr)
        if not re.fullmatch(r"[\d\s()+\-*]+", expr):
            raise ValueError("cannot evaluate %r" % expr)
        return int(eval(expr, {"__builtins__": {}}))

    def array(self, c_name):
        m = re.search(r"static const int8_t %s\[\]\s*=\s*(\w+);" % c_name, self.txt)
        body = self.defs[m.group(1)]
        return np.array([int(v) for v in body.strip("{} ").split(",") if v.strip()], dtype=np.int64).astype(np.int8)

    def weight(self, var):
        """nnom_weight_t / nnom_bias_t variable -> (int8 array, shift)"""
        m = re.search(r"static const nnom_(?:weight|bias)_t %s\s*=\s*\{\s*\(const void\*\)(\w+),\s*([^}]+)\}" % var, self.txt)
        return self.array(m.group(1)), self.value(m.group(2))


def split_args(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch == "," and depth == 0:
            out.append(cur.strip()); cur = ""
            continue
        depth += ch == "("
        depth -= ch == ")"
        cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def pair(call, name):
    m = re.fullmatch(r"%s\(\s*(\d+)\s*,\s*(\d+)\s*\)" % name, call)
    return int(m.group(1)), int(m.group(2))


def out_dim(n, k, s, same):
    return -(-n // s) if same else -(-(n - k + 1) // s)


def convert(src):
    H = Header(src)
    nodes, alias = [], {}                 # alias: header layer index -> node index

    def src_node(tok):
        return alias[int(re.fullmatch(r"layer\[(\d+)\]", tok).group(1))]

    for m in re.finditer(r"layer\[(\d+)\]\s*=\s*(.+?);\s*$", H.txt, re.M):
        idx, rhs = int(m.group(1)), m.group(2).strip()
        if rhs.startswith("Input("):
            h, w, c = [int(v) for v in re.match(r"Input\(shape\((\d+),\s*(\d+),\s*(\d+)\)", rhs).groups()]
            nodes.append(dict(op="input", ins=[], shape=(h, w, c)))
            alias[idx] = len(nodes) - 1
            continue
        call = re.fullmatch(r"model\.(hook|active|mergex|merge)\((.*)\)", rhs)
        if not call:
            raise ValueError("layer[%d]: cannot parse %r" % (idx, rhs))
        kind, args = call.group(1), split_args(call.group(2))
        if kind == "active":
            if not args[0].startswith("act_relu"):
                raise ValueError("layer[%d]: only act_relu tails are supported, got %s" % (idx, args[0]))
            n = src_node(args[1])
            if nodes[n]["op"] not in ("conv", "dense", "add"):
                raise ValueError("layer[%d]: ReLU tail on a %s node" % (idx, nodes[n]["op"]))
            if any(n in o["ins"] for o in nodes):
                raise ValueError("layer[%d]: activation attached after its layer already has readers" % idx)
            nodes[n]["relu"] = 1
            alias[idx] = n
            continue
        layer = args[0]
        ins = [src_node(a) for a in (args[2:] if kind == "mergex" else args[1:])]
        lname, largs = re.fullmatch(r"(\w+)\((.*)\)", layer).groups()
        largs = split_args(largs)
        ih, iw, ic = nodes[ins[0]]["shape"]
        if lname == "Conv2D":
            cout = int(largs[0]); kh, kw = pair(largs[1], "kernel"); sh, sw = pair(largs[2], "stride")
            same = int(largs[3] == "PADDING_SAME")
            w, rs = H.weight(largs[4].lstrip("&")); b, ls = H.weight(largs[5].lstrip("&"))
            assert w.size == cout * kh * kw * ic and b.size == cout, (idx, w.size)
            nodes.append(dict(op="conv", ins=ins, p=[cout, kh, kw, sh, sw, same, rs, ls, 0, ic], w=w, b=b, relu_at=8,
                              shape=(out_dim(ih, kh, sh, same), out_dim(iw, kw, sw, same), cout)))
        elif lname == "Dense":
            cout = int(largs[0])
            w, rs = H.weight(largs[1].lstrip("&")); b, ls = H.weight(largs[2].lstrip("&"))
            cin = ih * iw * ic
            assert w.size == cout * cin and b.size == cout, (idx, w.size, cout, cin)
            w = deinterleave_x4(w, cout, cin).reshape(-1)          # nnom_utils.py stores dense kernels x4-interleaved
            nodes.append(dict(op="dense", ins=ins, p=[cout, cin, rs, ls, 0], w=w, b=b, relu_at=4, shape=(1, 1, cout)))
        elif lname == "MaxPool":
            kh, kw = pair(largs[0], "kernel"); sh, sw = pair(largs[1], "stride")
            same = int(largs[2] == "PADDING_SAME")
            nodes.append(dict(op="maxpool", ins=ins, p=[kh, kw, sh, sw, same],
                              shape=(out_dim(ih, kh, sh, same), out_dim(iw, kw, sw, same), ic)))
        elif lname == "GlobalMaxPool":
            nodes.append(dict(op="gmaxpool", ins=ins, p=[], shape=(1, 1, ic)))
        elif lname == "Concat":
            assert int(largs[0]) in (-1, 3), "only channel-axis concat"
            shapes = [nodes[i]["shape"] for i in ins]
            if any(s != shapes[0] for s in shapes):
                raise ValueError("layer[%d]: this NNoM version sizes every concat input like the first one "
                                 "(nnom_concat.c:104, :208); unequal inputs are not reproducible" % idx)
            nodes.append(dict(op="concat", ins=ins, p=[], shape=(ih, iw, ic * len(ins))))
        elif lname == "Add":
            assert len(ins) == 2 and nodes[ins[0]]["shape"] == nodes[ins[1]]["shape"]
            nodes.append(dict(op="add", ins=ins, p=[H.value(largs[0]), 0], relu_at=1, shape=(ih, iw, ic)))
        elif lname == "Softmax":
            nodes.append(dict(op="softmax", ins=ins, p=[], shape=(ih, iw, ic)))
        elif lname == "Output":
            alias[idx] = ins[0]
            continue
        elif lname in ("Flatten", "Dropout"):
            alias[idx] = ins[0]
            continue
        else:
            raise ValueError("layer[%d]: %s is outside the supported vocabulary" % (idx, lname))
        alias[idx] = len(nodes) - 1
    for n in nodes:
        if n.get("relu"):
            n["p"][n["relu_at"]] = 1
    return H, nodes


def write(H, nodes, dst):
    out = bytearray(b"EDNNQ7\0\0") + struct.pack("<II", 2, len(nodes))
    for n in nodes:
        p = list(n["shape"]) if n["op"] == "input" else list(n["p"])
        ins = n["ins"] + [NONE] * (8 - len(n["ins"]))
        out += struct.pack("<II8I10I", OP[n["op"]], len(n["ins"]), *ins, *(p + [0] * (10 - len(p))))
        if "w" in n:
            out += n["w"].tobytes() + n["b"].tobytes()
            out += b"\0" * (-len(out) % 4)
    g = lambda k, d: H.value(k) if k in H.defs else d
    out += struct.pack("<iii", g("NNOM_INPUT_SCALE", 1), g("NNOM_INPUT_MIN", -128), g("NNOM_INPUT_MAX", 127))
    open(dst, "wb").write(out)
    return len(out)


def main():
    H, nodes = convert(sys.argv[1])
    for i, n in enumerate(nodes):
        print("%2d %-8s in=%-12s out=%s p=%s" % (i, n["op"], n["ins"], n["shape"], n.get("p", "")))
    print("wrote %s (%d bytes)" % (sys.argv[2], write(H, nodes, sys.argv[2])))


if __name__ == "__main__":
    main()
