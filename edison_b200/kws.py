"""Clip preparation and whole-clip inference as the reference's host script does it (audio/edison/kws/kws_on_mcu.py),
with the GPU library standing where the script talks to the MCU.

  fileInference  (kws_on_mcu.py:273-308): read WAV, cut / ZERO-pad to sample_len = 2 s = 32,000 samples, MFCC, inference
  frameInference (kws_on_mcu.py:310-345): same but EDGE-pad (np.pad mode='edge'), frames sent one by one to
                 appHifMfccAndInference (mfccAndInfereOnMCU :186-216), which returns the net input and the prediction
  float32 WAVs are scaled by 2**15 - 1 and truncated to int16 (:197-198, :326-327)

n_frames = 1 + (32,000 - 1024) // 1024 = 31 (audio/config.py:31): samples 31,744..31,999 of a 2 s clip are never used.
Nothing here computes features or the net: that is edison_b200.kws_batch (K1 + K2 through the C ABI).
"""
import numpy as np

from . import capi
from .hostif import read_wav_int16

FS = 16000                                        # audio/config.py:13
SAMPLE_LEN = int(2.0 * FS)                        # :12, :25-26
FRAME_LENGTH = FRAME_STEP = 1024                  # :20, :27
N_FRAMES = 1 + (SAMPLE_LEN - FRAME_LENGTH) // FRAME_STEP   # :31
NUM_MFCC = 13                                     # :24
assert N_FRAMES == capi.N_FRAMES and FRAME_LENGTH == capi.FRAME_LEN and NUM_MFCC == capi.N_COEFF


def to_int16(data):
    """float32 recordings -> int16 as mfccAndInfereOnMCU does (truncation toward zero, no clipping); int16 passes through"""
    data = np.asarray(data)
    if data.dtype == np.float32:
        return ((2 ** 15 - 1) * data).astype(np.int16)
    if data.dtype != np.int16:
        raise TypeError("expected int16 or float32 samples, got %s" % data.dtype)
    return data


def fit_clip(data, pad="zero"):
    """cut or pad one recording to SAMPLE_LEN samples; pad='zero' (fileInference) or 'edge' (frameInference)"""
    if pad not in ("zero", "edge"):
        raise ValueError("pad must be 'zero' or 'edge'")
    data = to_int16(data).reshape(-1)
    if data.size >= SAMPLE_LEN:
        return np.ascontiguousarray(data[:SAMPLE_LEN])
    if data.size == 0 and pad == "edge":
        raise ValueError("cannot edge-pad an empty recording")          # np.pad(mode='edge') raises on empty input too
    return np.pad(data, (0, SAMPLE_LEN - data.size), mode="constant" if pad == "zero" else "edge")


def infer_clips(clips, want_softmax=True):
    """clips: int16 [n, >= 31,744] (rows as produced by fit_clip) -> (net_in int8 [n, 31, 13], logits int8 [n, 10], softmax | None)
    -- what appHifMfccAndInference returns for each clip"""
    clips = np.ascontiguousarray(clips, dtype=np.int16)
    if clips.ndim == 1:
        clips = clips.reshape(1, -1)
    if clips.shape[1] < N_FRAMES * FRAME_LENGTH:
        raise ValueError("a clip needs %d samples, got %d" % (N_FRAMES * FRAME_LENGTH, clips.shape[1]))
    lg, sm, ni = capi.kws_batch(clips, clips.shape[0], clip_stride=clips.shape[1], want_softmax=want_softmax, want_net_in=True)
    return ni.reshape(-1, N_FRAMES, NUM_MFCC), lg, sm


def file_inference(path_or_samples, pad="zero"):
    """one recording (WAV path or sample array) -> (net_in [31, 13], logits [10], softmax [10], class index)"""
    data = read_wav_int16(path_or_samples, expect_rate=FS) if isinstance(path_or_samples, str) else path_or_samples
    ni, lg, sm = infer_clips(fit_clip(data, pad))
    return ni[0], lg[0], sm[0], int(sm[0].argmax())


def frame_inference(path_or_samples):
    """frameInference's variant: edge padding"""
    return file_inference(path_or_samples, pad="edge")
