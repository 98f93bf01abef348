"""TEST INFRASTRUCTURE -- CPU restatement (numpy, float64 accumulation) of the reference's float CNN.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product never does.

PARITY UNPINNED at the keras boundary: the reference runs this model through ST's closed X-CUBE-AI 5.0.0 runtime
(firmware/src/ai/cube, libNetworkRuntime*.a for Cortex-M4) or keras 2.3.1 / TensorFlow 1.15 -- neither can run here, and the
reference stores no input/output tensors for it (its own checks are statistical: README.md:118-127).  This restatement follows
  topology      audio/edison/train/kws_keras.py:371-397 (Conv 16@5x5, 32@3x3, 64@3x3, 32@3x3, all 'valid', stride 1;
                BatchNorm folded into the conv weights by Cube.AI; ReLU; MaxPool (2,1) after the first two; Dense 10; Softmax)
  layer order   firmware/src/ai/cube/kws/kws.c:306-420 (conv2d_nl_pool = conv -> ReLU -> max pool (1,2) in Cube's (x,y) order)
  tensors       kws.c:213-291 (weights O,H,W,I; dense [out][in]; activations H,W,C; flatten of (1,3,32) in H,W,C order)
  weights       kws.c:488-516 byte offsets into kws_data.c (converted by tools/convert_cube_weights.py)
  input         float [31][13][1] = MFCC values cast to float, unscaled (firmware/src/app.c:679-683)
  softmax       natural e (forward_sm)."""
import struct

import numpy as np


def load(path):
    raw = open(path, "rb").read()
    assert raw[:8] == b"EDCUBEF\0"
    ver, n = struct.unpack_from("<II", raw, 8)
    assert ver == 1
    off, layers = 16, []
    for _ in range(n):
        kind, co, kh, kw, ci, relu, pool = struct.unpack_from("<7I", raw, off)
        off += 28
        w = np.frombuffer(raw, dtype="<f4", count=co * kh * kw * ci, offset=off).reshape(co, kh, kw, ci)
        off += w.size * 4
        b = np.frombuffer(raw, dtype="<f4", count=co, offset=off)
        off += co * 4
        layers.append(dict(kind=kind, w=w, b=b, relu=relu, pool=pool))
    assert off == len(raw)
    return layers


def forward(layers, x, dtype=np.float64):
    """x: [n][31][13] (or [n][403]) -> (logits [n][10], softmax [n][10]) computed in `dtype`."""
    a = np.asarray(x, dtype=dtype).reshape(-1, 31, 13, 1)
    for L in layers:
        w, b = L["w"].astype(dtype), L["b"].astype(dtype)
        if L["kind"] == 1:
            co, kh, kw, ci = w.shape
            n, h, wd, _ = a.shape
            ho, wo = h - kh + 1, wd - kw + 1
            out = np.zeros((n, ho, wo, co), dtype=dtype)
            for ky in range(kh):
                for kx in range(kw):
                    out += np.einsum("nhwc,oc->nhwo", a[:, ky:ky + ho, kx:kx + wo, :], w[:, ky, kx, :])
            out += b
            if L["relu"]:
                out = np.maximum(out, 0)
            if L["pool"]:
                hp = ho // 2
                out = np.maximum(out[:, 0:2 * hp:2], out[:, 1:2 * hp:2])
            a = out
        else:
            a = a.reshape(a.shape[0], -1) @ w.reshape(w.shape[0], -1).T + b
    logits = a
    e = np.exp(logits - logits.max(axis=1, keepdims=True))
    return logits, e / e.sum(axis=1, keepdims=True)
