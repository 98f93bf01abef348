"""Index-flow prototype of K3 (float MFCC): 512-point complex FFT as three radix-8 DIF passes with the lane/register
assignment and shared-memory layouts of mfcc_f32_kernel.cu, then the real-input split.  Checks against numpy and
counts shared-memory bank conflicts (64-bit accesses, half-warp phases)."""
import numpy as np

rng = np.random.default_rng(1)
x = rng.integers(-32768, 32767, 1024).astype(np.float64)
z = x[0::2] + 1j * x[1::2]
L = np.arange(32)
W = lambda N, e: np.exp(-2j * np.pi * (e % N) / N)


def conflicts(addrs):
    """addrs: int array [32] of float2 indices; returns wavefronts needed (ideal 2 for 64-bit)"""
    w = 0
    for half in (addrs[:16], addrs[16:]):
        banks = {}
        for a in set(half.tolist()):
            banks.setdefault(a % 16, set()).add(a)
        w += max(len(v) for v in banks.values())
    return w


def r8(v):  # v[8] -> DFT8 (DIF), arrays over lanes
    a = [v[i] + v[i + 4] for i in range(4)]
    b = [v[i] - v[i + 4] for i in range(4)]
    b[1] = b[1] * W(8, 1); b[2] = b[2] * -1j; b[3] = b[3] * W(8, 3)
    out = [None] * 8
    for c, base in ((a, 0), (b, 1)):
        e0, e1, e2, e3 = c[0] + c[2], c[0] - c[2], c[1] + c[3], (c[1] - c[3]) * -1j
        out[base + 0], out[base + 4], out[base + 2], out[base + 6] = e0 + e2, e0 - e2, e1 + e3, e1 - e3
    return out


worst = 0
buf = np.zeros(600, dtype=complex)
# pass 1: lane l, h: m = l + 32h, inputs n = 64 n2 + m
y1 = {}
for h in range(2):
    m = L + 32 * h
    o = r8([z[64 * n2 + m] for n2 in range(8)])
    for r in range(8):
        y1[(r, h)] = o[r] * W(512, m * r)
# exchange 1: addr = 72 r + m
for (r, h), v in y1.items():
    a = 72 * r + L + 32 * h
    worst = max(worst, conflicts(a)); buf[a] = v
y2 = {}
for h in range(2):
    r = (L >> 3) + 4 * h; m0 = L & 7
    v = []
    for m1 in range(8):
        a = 72 * r + 8 * m1 + m0
        worst = max(worst, conflicts(a)); v.append(buf[a].copy())
    o = r8(v)
    for rp in range(8):
        y2[(rp, h)] = o[rp] * W(64, m0 * rp)
# exchange 2: addr = 72 r' + 8 r + (m0 ^ r)
buf2 = np.zeros(600, dtype=complex)
for (rp, h), v in y2.items():
    r = (L >> 3) + 4 * h; m0 = L & 7
    a = 72 * rp + 8 * r + (m0 ^ r)
    worst = max(worst, conflicts(a)); buf2[a] = v
Z = np.zeros(513, dtype=complex)
for h in range(2):
    r = L & 7; rp = (L >> 3) + 4 * h
    v = []
    for m0 in range(8):
        a = 72 * rp + 8 * r + (m0 ^ r)
        worst = max(worst, conflicts(a)); v.append(buf2[a].copy())
    o = r8(v)
    for q in range(8):
        k = 64 * q + 8 * rp + r          # = 64 q + l + 32 h
        assert np.array_equal(k, 64 * q + L + 32 * h)
        worst = max(worst, conflicts(k)); Z[k] = o[q]
ref = np.fft.fft(z)
print("fft512 err", np.abs(Z[:512] - ref).max() / np.abs(ref).max(), "worst wavefronts", worst)
# real split: pairs (k, 512-k), k = l + 32 j, j < 8
Z[512] = Z[0]
X = np.zeros(513, dtype=complex)
for j in range(8):
    k = L + 32 * j
    A, B = Z[k], np.conj(Z[512 - k])
    xe, xo = (A + B) / 2, -1j * (A - B) / 2
    t = W(1024, k) * xo
    X[k] = xe + t
    X[512 - k] = np.conj(xe - t)
X[256] = np.conj(Z[256])
full = np.fft.fft(x)
print("rfft err", np.abs(X - full[:513]).max() / np.abs(full).max())
