#!/usr/bin/env python3
"""Convert the firmware's generated ``mel_constants.h`` into the flat blob libedison_b200 loads.

The reference table is firmware/src/audio/mel_constants.h (written by
audio/edison/mfcc/mfcc_on_mcu.py:68-145 from gen_mel_weight_matrix): ``melMtxCompact[915]`` (:2586),
``melCompFStarts[32]`` (:2645), ``melCompFCount[32]`` (:2648), ``MEL_MTX_SCALE 128`` (:7).
The blob is MODEL DATA (like kws_nnom_q7.bin), not source.

Blob layout (little endian):
  char[8]  "EDMELQ15"
  u32 version (1), u32 n_mel, u32 n_weights, u32 scale
  i16 starts[n_mel], i16 counts[n_mel], i16 weights[n_weights], zero padding to 4 bytes

usage: convert_mel_constants.py <mel_constants.h> <out.bin>
"""
import re
import struct
import sys

import numpy as np


def array(txt, name):
    m = re.search(r"%s\[(\d+)\]\s*=\s*\{([^}]*)\}" % name, txt, re.S)
    vals = np.array([int(v) for v in m.group(2).replace("\n", " ").split(",") if v.strip()], dtype=np.int16)
    assert vals.size == int(m.group(1)), (name, vals.size)
    return vals


def main():
    src, dst = sys.argv[1:3]
    txt = open(src).read()
    scale = int(re.search(r"#define\s+MEL_MTX_SCALE\s+(\d+)", txt).group(1))
    w = array(txt, "melMtxCompact")
    starts = array(txt, "melCompFStarts")
    counts = array(txt, "melCompFCount")
    assert starts.size == counts.size and counts.sum() == w.size
    out = bytearray(b"EDMELQ15")
    out += struct.pack("<IIII", 1, starts.size, w.size, scale)
    out += starts.tobytes() + counts.tobytes() + w.tobytes()
    out += b"\0" * (-len(out) % 4)
    open(dst, "wb").write(out)
    print("wrote %s: %d filters, %d weights, bins %d..%d" % (dst, starts.size, w.size, starts.min(), (starts + counts).max() - 1))


if __name__ == "__main__":
    main()
