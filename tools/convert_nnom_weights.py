#!/usr/bin/env python3
"""Convert an NNoM-generated ``weights.h`` (keras -> NNoM, nnom_utils.py) into the flat
binary blob that libedison_b200 and the oracle load.

The reference model is firmware/src/ai/nnom/kws_nnom/weights.h (graph at :138-161, shift
macros :5-105).  The blob is MODEL DATA (like a checkpoint), not source: int8 weights in the
reference's own O,H,W,I order, int8 biases, and the two per-layer shifts
(OUTPUT_RSHIFT = in_shift + kernel_shift - out_shift, BIAS_LSHIFT = in_shift + kernel_shift -
bias_shift; weights.h:72-105).  The dense kernel is stored by NNoM in the CMSIS-NN "x4
interleaved" order (arm_fully_connected_q7_opt.c:129-185, nnom_local.c:804-921); it is
de-interleaved here into plain row-major [out][in].

Blob layout (little endian):
  char[8]  "EDNNQ7\\0\\0"
  u32      version (1), u32 n_layers
  per layer: u32 kind (1 conv, 2 dense), u32 cout, u32 kh, u32 kw, u32 cin,
             u32 out_rshift, u32 bias_lshift, u32 relu, u32 pool_h (0 = none),
             int8 w[cout*kh*kw*cin], int8 b[cout], zero padding to 4 bytes
  trailer: i32 input_scale, i32 input_min, i32 input_max, u32 in_h, u32 in_w, u32 in_c

usage: convert_nnom_weights.py <weights.h> <out.bin>
"""
import re
import struct
import sys

import numpy as np


def parse(path):
    txt = open(path).read()
    arrays = {m.group(1): np.array([int(v) for v in m.group(2).split(",") if v.strip()], dtype=np.int64)
              for m in re.finditer(r"#define\s+(\w+)\s+\{([^}]*)\}", txt)}
    ints = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define\s+(\w+)\s+\(?(-?\d+)\)?\s*$", txt, re.M)}
    return txt, arrays, ints


def deinterleave_x4(w, rows, cols):
    """Invert convert_to_x4_q7_weights: read order of local_fully_connected_q7_opt (nnom_local.c:831-874)."""
    out = np.zeros((rows, cols), dtype=np.int8)
    p = 0
    assert cols % 4 == 0
    for r0 in range(0, rows - rows % 4, 4):
        for c0 in range(0, cols, 4):
            blk = w[p:p + 16]
            p += 16
            # (row offset, col offset) of the 16 consecutive bytes
            order = [(0, 0), (1, 0), (0, 2), (1, 2), (2, 0), (3, 0), (2, 2), (3, 2),
                     (0, 1), (1, 1), (0, 3), (1, 3), (2, 1), (3, 1), (2, 3), (3, 3)]
            for v, (dr, dc) in zip(blk, order):
                out[r0 + dr, c0 + dc] = v
    for r in range(rows - rows % 4, rows):
        out[r, :] = w[p:p + cols]
        p += cols
    assert p == w.size
    return out


def main():
    src, dst = sys.argv[1:3]
    txt, arrays, ints = parse(src)
    # graph of the reference model, read from nnom_model_create (weights.h:138-161)
    convs = re.findall(r"Conv2D\((\d+),\s*kernel\((\d+),\s*(\d+)\).*?&(\w+)_w,\s*&\w+_b\)", txt)
    dense = re.findall(r"Dense\((\d+),\s*&(\w+)_w", txt)
    pools = re.findall(r"layer\[(\d+)\] = model\.hook\(MaxPool\(kernel\((\d+),\s*(\d+)\),\s*stride\((\d+),\s*(\d+)\)", txt)
    in_shape = [int(v) for v in re.search(r"Input\(shape\((\d+),\s*(\d+),\s*(\d+)\)", txt).groups()]
    # which conv is followed by a pool: walk the layer list in order
    order = re.findall(r"layer\[\d+\] = model\.(?:hook|active)\((\w+)\(", txt)
    layers = []
    cin = in_shape[2]
    ci = 0
    prev_shift_name = "INPUT_1"
    seq = [o for o in order]
    idx = 0
    for k, name in enumerate(seq):
        if name == "Conv2D":
            cout, kh, kw, lname = convs[ci]
            ci += 1
            cout, kh, kw = int(cout), int(kh), int(kw)
            up = lname.upper()
            w = arrays[up + "_KERNEL_0"].astype(np.int8)
            b = arrays[up + "_BIAS_0"].astype(np.int8)
            assert w.size == cout * kh * kw * cin, (lname, w.size)
            relu = int(k + 1 < len(seq) and seq[k + 1] == "act_relu")
            pool = 0
            j = k + 1 + relu
            if j < len(seq) and seq[j] == "MaxPool":
                ph, pw, sh, sw = [int(v) for v in pools[sum(1 for s in seq[:j] if s == "MaxPool")][1:]]
                assert (pw, sw) == (1, 1) and ph == sh, "only (n,1)/(n,1) pooling is on the reference path"
                pool = ph
            layers.append(dict(kind=1, cout=cout, kh=kh, kw=kw, cin=cin, w=w, b=b, relu=relu, pool=pool,
                               rshift=None, lshift=None, name=up))
            cin = cout
        elif name == "Dense":
            cout, lname = dense[0]
            cout = int(cout)
            up = lname.upper()
            w = arrays[up + "_KERNEL_0"].astype(np.int8)
            b = arrays[up + "_BIAS_0"].astype(np.int8)
            n_in = w.size // cout
            w = deinterleave_x4(w, cout, n_in).reshape(-1)
            layers.append(dict(kind=2, cout=cout, kh=1, kw=1, cin=n_in, w=w, b=b, relu=0, pool=0, name=up))
    # shifts: evaluate the macros the way the header does (weights.h:72-105)
    for L in layers:
        m = re.search(r"#define %s_OUTPUT_RSHIFT \((\w+)\+(\w+)-(\w+)\)" % L["name"], txt)
        L["rshift"] = ints[m.group(1)] + ints[m.group(2)] - ints[m.group(3)]
        m = re.search(r"#define %s_BIAS_LSHIFT\s+\((\w+)\+(\w+)-(\w+)\)" % L["name"], txt)
        L["lshift"] = ints[m.group(1)] + ints[m.group(2)] - ints[m.group(3)]
        assert L["rshift"] >= 1 and L["lshift"] >= 0
    out = bytearray(b"EDNNQ7\0\0")
    out += struct.pack("<II", 1, len(layers))
    for L in layers:
        out += struct.pack("<9I", L["kind"], L["cout"], L["kh"], L["kw"], L["cin"], L["rshift"], L["lshift"], L["relu"], L["pool"])
        out += L["w"].tobytes() + L["b"].tobytes()
        out += b"\0" * (-len(out) % 4)
    # the NNOM_INPUT_* macros are an addition of the reference's own generator; upstream NNoM headers lack them
    out += struct.pack("<iiiIII", ints.get("NNOM_INPUT_SCALE", 1), ints.get("NNOM_INPUT_MIN", -128), ints.get("NNOM_INPUT_MAX", 127),
                       *in_shape)
    open(dst, "wb").write(out)
    for L in layers:
        print("%-10s kind=%d cout=%d k=%dx%d cin=%d rshift=%d lshift=%d relu=%d pool=%d" %
              (L["name"], L["kind"], L["cout"], L["kh"], L["kw"], L["cin"], L["rshift"], L["lshift"], L["relu"], L["pool"]))
    print("wrote %s (%d bytes)" % (dst, len(out)))


if __name__ == "__main__":
    main()
