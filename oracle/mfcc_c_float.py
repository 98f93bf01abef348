"""TEST INFRASTRUCTURE -- numpy restatement of the reference's float MFCC in C (firmware/src/audio/mfcc.c).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product never does.

Follows mfcc.c: mfcc_create :47-83 (Hann window :67-68, padded length :60), create_dct_matrix :101-115,
create_mel_fbank :117-171 (26 bins 20..4000 Hz, mfcc.h:28-31, float32 MelScale mfcc.h:53-55), mfcc_compute :173-255
(x[0] left unscaled :178-179, pre-emphasis :180-184, window, real FFT, power, sqrt(power) * weight :219-220,
FLT_MIN for empty bands :225-227, logf, DCT rows offset..num-1, round(sum * 2^dec_bits) saturated to [-128, 127]).
float32 is kept for the table construction (it decides which bins belong to a filter); the FFT runs in float64.
Pinned by tests/golden/mfcc_c_float.npz = outputs of the unmodified mfcc.c built for the host (oracle/_ref/libref_mfcc_f32.so):
agreement is exact except where round() sits within float32 noise of a tie (<= 1 LSB, counted in the test)."""
import numpy as np

f32 = np.float32
SAMP_FREQ, N_FBANK, MEL_LO, MEL_HI = 16000, 26, 20, 4000


def mel_scale(f):
    return f32(1127.0) * np.log(f32(1.0) + np.asarray(f, dtype=f32) / f32(700.0), dtype=f32)


class MfccC:
    def __init__(self, num_mfcc=13, offset=1, frame_len=512, dec_bits=8, preemph=0.97):
        self.num, self.offset, self.frame_len, self.dec_bits, self.pre = num_mfcc, offset, frame_len, dec_bits, f32(preemph)
        self.padded = int(2 ** int(np.ceil(np.log2(frame_len))))
        i = np.arange(frame_len, dtype=f32)
        self.window = (f32(0.5) - f32(0.5) * np.cos(f32(6.283185307179586) * i / f32(frame_len), dtype=f32)).astype(f32)
        nb = self.padded // 2
        width = f32(SAMP_FREQ) / f32(self.padded)
        lo, hi = mel_scale(MEL_LO), mel_scale(MEL_HI)
        delta = f32((hi - lo) / f32(N_FBANK + 1))
        mel = mel_scale(width * np.arange(nb, dtype=f32))
        self.first, self.last, self.w = [], [], []
        for b in range(N_FBANK):
            left, center, right = f32(lo + f32(b) * delta), f32(lo + f32(b + 1) * delta), f32(lo + f32(b + 2) * delta)
            inside = (mel > left) & (mel < right)
            idx = np.nonzero(inside)[0]
            w = np.where(mel <= center, (mel - left) / f32(center - left), (right - mel) / f32(right - center)).astype(f32)
            self.first.append(int(idx[0])); self.last.append(int(idx[-1]))
            self.w.append(np.where(inside, w, f32(0))[idx[0]:idx[-1] + 1])
        norm = np.sqrt(f32(2.0) / f32(N_FBANK), dtype=f32)
        n = np.arange(N_FBANK, dtype=f32)
        self.dct = np.array([norm * np.cos(f32(np.pi) / f32(N_FBANK) * (n + f32(0.5)) * f32(k), dtype=f32) for k in range(num_mfcc)], dtype=f32)

    def compute(self, audio):
        """audio: int16[frame_len] -> (int8[num - offset], float sums before quantisation)"""
        a = np.asarray(audio[:self.frame_len], dtype=f32)
        fr = np.zeros(self.padded, dtype=f32)
        fr[0] = a[0]
        fr[1:self.frame_len] = (a[1:] - a[:-1] * self.pre) / f32(32768)
        fr[:self.frame_len] *= self.window
        spec = np.fft.rfft(fr.astype(np.float64))
        power = (spec.real ** 2 + spec.imag ** 2).astype(f32)
        mag = np.sqrt(power, dtype=f32)
        mel = np.zeros(N_FBANK, dtype=f32)
        for b in range(N_FBANK):
            e = f32(0)
            for v, w in zip(mag[self.first[b]:self.last[b] + 1], self.w[b]):
                e = f32(e + v * w)
            mel[b] = e if e != 0 else np.finfo(f32).tiny
        lm = np.log(mel, dtype=f32)
        sums = np.array([np.sum(self.dct[k].astype(np.float64) * lm) for k in range(self.offset, self.num)]) * (1 << self.dec_bits)
        q = np.clip(np.sign(sums) * np.floor(np.abs(sums) + 0.5), -128, 127).astype(np.int8)   # C round(): half away from zero
        return q, sums

    def batch(self, pcm, n_frames, hop):
        out = np.zeros((n_frames, self.num - self.offset), dtype=np.int8)
        sums = np.zeros((n_frames, self.num - self.offset))
        for f in range(n_frames):
            out[f], sums[f] = self.compute(pcm[f * hop:f * hop + self.frame_len])
        return out, sums
