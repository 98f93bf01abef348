#!/usr/bin/env python3
"""Convert the Cube.AI weight array of the float model (firmware/src/ai/cube/kws/kws_data.c, 171,944 bytes of little
endian float32) into the blob libedison_b200 and the oracle load.  MODEL DATA (like a checkpoint), not source.

The tensor layout is the one kws.c declares: byte offsets kws.c:488-516, shapes/strides :213-291
(conv weights O,H,W,I; dense [out][in]; BatchNorm already folded by X-CUBE-AI 5.0.0).

Blob layout (little endian):
  char[8]  "EDCUBEF\\0"
  u32      version (1), u32 n_layers (5)
  per layer: u32 kind (1 conv, 2 dense), u32 cout, u32 kh, u32 kw, u32 cin, u32 relu, u32 pool_h (0 = none),
             f32 w[cout*kh*kw*cin], f32 b[cout]

usage: convert_cube_weights.py <kws_data.c> <out.bin>
"""
import re
import struct
import sys

import numpy as np

# (kind, cout, kh, kw, cin, relu, pool_h, weight byte offset, bias byte offset)   kws.c:488-516, :306-420
LAYERS = [
    (1, 16, 5, 5, 1, 1, 2, 0, 1600),
    (1, 32, 3, 3, 16, 1, 2, 1664, 20096),
    (1, 64, 3, 3, 32, 1, 0, 20224, 93952),
    (1, 32, 3, 3, 64, 1, 0, 94208, 167936),
    (2, 10, 1, 1, 96, 0, 0, 168064, 171904),
]


def main(src, dst):
    txt = open(src).read()
    m = re.search(r"s_kws_weights\[\s*(\d+)\s*\]\s*=\s*\{([^}]*)\}", txt)
    n = int(m.group(1))
    raw = bytes(int(v, 16) for v in re.findall(r"0x[0-9a-fA-F]{2}", m.group(2)))
    assert len(raw) == n == 171944, (len(raw), n)
    out = bytearray(b"EDCUBEF\0" + struct.pack("<II", 1, len(LAYERS)))
    for kind, co, kh, kw, ci, relu, pool, wo, bo in LAYERS:
        nw = co * kh * kw * ci
        w = np.frombuffer(raw, dtype="<f4", count=nw, offset=wo)
        b = np.frombuffer(raw, dtype="<f4", count=co, offset=bo)
        assert np.isfinite(w).all() and np.isfinite(b).all()
        assert wo + 4 * nw <= bo
        out += struct.pack("<7I", kind, co, kh, kw, ci, relu, pool) + w.tobytes() + b.tobytes()
    open(dst, "wb").write(out)
    print("wrote %s: %d bytes" % (dst, len(out)))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
