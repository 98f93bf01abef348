"""TEST INFRASTRUCTURE -- CPU restatement (numpy, float64) of the reference's host-side float MFCCs.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the product never does.

Follows audio/edison/mfcc/mfcc_utils.py of the reference:
  hertz_to_mel            :30-34     gen_mel_weight_matrix :36-73     frames :16-27
  mfcc  (host float path) :134-199   mfcc_mcu (float imitation of the firmware path) :255-323
with the constants of audio/config.py:12-36 (fs 16000, frame 1024, hop 1024, 32 mel bins 80..7600 Hz, scale 128).
Pinned by tests/golden/mfcc_float.npz, which holds outputs of the reference module itself
(tests/golden/make_golden_float.py, run in the build container where /root/reference is mounted)."""
import numpy as np
from scipy.fftpack import dct

FS, FRAME, N_MEL, F_LO, F_HI, MEL_SCALE = 16000, 1024, 32, 80.0, 7600.0, 128
MEL_Q, MEL_BREAK = 1127.0, 700.0      # config.py:33-34


def hertz_to_mel(f):
    return MEL_Q * np.log(1.0 + (f / MEL_BREAK))


def mel_weight_matrix(n_bins, n_mel=N_MEL, fs=FS, lo=F_LO, hi=F_HI):
    """[n_bins][n_mel]; bin frequencies are linspace(0, fs/2, n_bins) with the DC row zeroed (mfcc_utils.py:43-73).
    Note n_bins = 512 for mfcc() (the spectrum is cut to frame_len//2 bins, so the bin spacing is 8000/511 Hz, not
    fs/1024 -- kept as the reference has it) and 513 for mfcc_mcu()."""
    lin = np.linspace(0, fs / 2.0, n_bins)[1:]
    bins_mel = hertz_to_mel(lin)[:, None]
    edges = np.linspace(hertz_to_mel(lo), hertz_to_mel(hi), n_mel + 2)
    lower, center, upper = edges[None, :-2], edges[None, 1:-1], edges[None, 2:]
    w = np.maximum(0, np.minimum((bins_mel - lower) / (center - lower), (upper - bins_mel) / (upper - center)))
    return np.pad(w, [[1, 0], [0, 0]])


def n_frames(n_samples, frame_len=FRAME, frame_step=FRAME):
    return 1 + (n_samples - frame_len) // frame_step


def mfcc_host(data):
    """mfcc_utils.mfcc: |FFT| first 512 bins -> mel -> log(. + 1e-6) -> DCT-II / sqrt(64).  Returns float64 [frames][32]."""
    data = np.asarray(data)
    w = mel_weight_matrix(FRAME // 2)
    out = np.zeros((n_frames(data.shape[0]), N_MEL))
    for i in range(out.shape[0]):
        spec = np.abs(np.fft.fft(data[i * FRAME:(i + 1) * FRAME])[:FRAME // 2])
        out[i] = dct(np.log(np.dot(spec, w) + 1e-6), type=2) / np.sqrt(2 * N_MEL)
    return out


def mfcc_mcu(data, use_log=False):
    """mfcc_utils.mfcc_mcu: fft/1024, |.|/sqrt(2), 513 bins, weights 128 W then /128, optional log, DCT-II / 64."""
    data = np.asarray(data)
    w = MEL_SCALE * mel_weight_matrix(FRAME // 2 + 1)
    out = np.zeros((n_frames(data.shape[0]), N_MEL))
    for i in range(out.shape[0]):
        chunk = np.array(data[i * FRAME:(i + 1) * FRAME])
        spec = 1.0 / np.sqrt(2) * np.abs(1.0 / 1024 * np.fft.fft(chunk))
        mel = np.dot(spec[:FRAME // 2 + 1], w)
        mel /= MEL_SCALE
        if use_log:
            mel = np.log(mel + 1e-6)
        out[i] = 1.0 / 64 * dct(mel, type=2)
    return out
