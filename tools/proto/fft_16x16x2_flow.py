#!/usr/bin/env python3
"""Index flow of K3's 512-point complex FFT as 16 x 16 x 2 (two shared-memory exchanges instead of three) plus the real-input split that
absorbs the last radix-2 stage, checked against numpy.fft.rfft.  Mirrors the lane / register assignment of mfcc_float.cu:
  pass A: lane m (0..31) holds z[32 a + m], a = 0..15            -> Y1[k0][m] = sum_a z[32a+m] W16^{a k0},  times W256^{(m >> 1) k0}  (m = 2 b + c)
  exchange 1: lane m writes xch[k0 * 34 + m]; lane (k0, c) = (lane >> 1, lane & 1) reads xch[k0 * 34 + 2 b + c], b = 0..15
  pass B: Y2[k0, k1; c] = sum_b Y1'[k0][2b + c] W16^{b k1}
  exchange 2: lane (k0, c) writes y2[(k0 + 16 k1) * 2 + c]
  split: for k in 0..127: P, Q = y2[2k], y2[2k+1]; t = W512^k Q; Z[k] = P + t, Z[k+256] = P - t; likewise k' = 256 - k; then the usual real split on
         the pairs (k, 512 - k) and (256 - k, 256 + k)."""
import numpy as np

def main():
    rng = np.random.default_rng(1)
    x = rng.integers(-32768, 32768, 1024).astype(np.float64)
    z = x[0::2] + 1j * x[1::2]
    W = lambda N, e: np.exp(-2j * np.pi * e / N)
    xch = np.zeros(16 * 34, complex)
    for m in range(32):                       # pass A
        v = np.array([z[32 * a + m] for a in range(16)])
        for k0 in range(16):
            y = sum(v[a] * W(16, a * k0) for a in range(16)) * W(256, (m >> 1) * k0)
            xch[k0 * 34 + m] = y
    y2 = np.zeros(512, complex)
    for lane in range(32):                    # pass B
        k0, c = lane >> 1, lane & 1
        v = np.array([xch[k0 * 34 + 2 * b + c] for b in range(16)])
        for k1 in range(16):
            y2[(k0 + 16 * k1) * 2 + c] = sum(v[b] * W(16, b * k1) for b in range(16))
    mag = np.zeros(513)
    def pair(a, bc, w):                       # the kernel's split: returns 2 X[k], 2 X[512 - k]
        er, ei = a.real + bc.real, a.imag - bc.imag
        dr, di = a.real - bc.real, a.imag + bc.imag
        tr, ti = w.real * di + w.imag * dr, w.imag * di - w.real * dr
        return complex(er + tr, ei + ti), complex(er - tr, ei - ti)
    for k in range(0, 129):
        P, Q = y2[2 * k], y2[2 * k + 1]
        w1024 = W(1024, k)
        w512 = w1024 * w1024
        t = w512 * Q
        Zk, Zk256 = P + t, P - t
        if k == 0:
            x1, x2 = pair(Zk, Zk, w1024)
            mag[0], mag[512] = abs(x1), abs(x2)
            mag[256] = 2 * abs(Zk256)
            continue
        if k == 128:
            x1, x2 = pair(Zk, Zk256, w1024)
            mag[128], mag[384] = abs(x1), abs(x2)
            continue
        kp = 256 - k
        Pp, Qp = y2[2 * kp], y2[2 * kp + 1]
        tp = -np.conj(w512) * Qp              # W512^{256 - k} = -conj(W512^k)
        Zkp, Z512k = Pp + tp, Pp - tp         # Z[256 - k], Z[512 - k]
        x1, x2 = pair(Zk, Z512k, w1024)
        mag[k], mag[512 - k] = abs(x1), abs(x2)
        wp = -1j * np.conj(w1024)             # W1024^{256 - k}
        x1, x2 = pair(Zkp, Zk256, wp)
        mag[kp], mag[256 + k] = abs(x1), abs(x2)
    ref = 2 * np.abs(np.fft.rfft(x))
    err = np.max(np.abs(mag - ref) / (1 + ref))
    print("max relative error vs numpy rfft:", err)
    assert err < 1e-12
    # bank check of exchange 1's reads (64-bit units, 16 lanes per wavefront, 16 bank pairs)
    for b in range(16):
        for half in range(2):
            banks = {(((l >> 1) * 34 + 2 * b + (l & 1)) % 16) for l in range(16 * half, 16 * half + 16)}
            assert len(banks) == 16
    print("exchange 1 reads: conflict-free with row stride 34")

if __name__ == "__main__":
    main()
