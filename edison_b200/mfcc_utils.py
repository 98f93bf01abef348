"""Host-side mirror of the reference's audio/edison/mfcc/mfcc_utils.py for the float MFCC path: same function names and
argument lists (mfcc :134, mfcc_mcu :255), computed by K3 (csrc/mfcc_float.cu, fp64 instantiation) on the GPU.  Only the reference's own
configuration (audio/config.py:12-36) has kernels; anything else raises.  The per-frame dicts carry 't_start', 't_end'
and 'mfcc' (the reference also stores its intermediates, which stay on the device here)."""
import numpy as np

from . import capi

_CFG = dict(fs=16000, frame_len=1024, frame_step=1024, fft_len=1024, mel_nbins=32, mel_lower_hz=80.0, mel_upper_hz=7600.0)


def _check(**kw):
    for k, v in kw.items():
        if float(v) != float(_CFG[k]):
            raise ValueError("edison_b200 float MFCC is built for %s=%s (audio/config.py), got %s" % (k, _CFG[k], v))


def _frames(data, nSamples, frame_count, mode):
    data = np.asarray(data)
    if data.dtype != np.int16:
        if not np.array_equal(data, np.round(data)) or np.abs(data).max(initial=0) > 32768:
            raise ValueError("the GPU path takes the int16 PCM the reference reads from WAV files")
        data = data.astype(np.int16)
    if frame_count == 0:
        frame_count = 1 + (int(nSamples) - 1024) // 1024
    co = capi.mfcc_float_clips(data, frame_count, 1, 1024, mode, np.float64)
    return [{"t_start": i * 1024 / 16000, "t_end": (i * 1024 + 1024) / 16000, "mfcc": co[i]} for i in range(frame_count)]


def mfcc(data, fs, nSamples, frame_len, frame_step, frame_count, fft_len, mel_nbins, mel_lower_hz, mel_upper_hz, dummy=None):
    _check(fs=fs, frame_len=frame_len, frame_step=frame_step, fft_len=fft_len, mel_nbins=mel_nbins,
           mel_lower_hz=mel_lower_hz, mel_upper_hz=mel_upper_hz)
    return _frames(data, nSamples, frame_count, capi.MFCC_F32_HOST)


def mfcc_mcu(data, fs, nSamples, frame_len, frame_step, frame_count, fft_len, mel_nbins, mel_lower_hz, mel_upper_hz,
             mel_mtx_scale, use_log=False):
    _check(fs=fs, frame_len=frame_len, frame_step=frame_step, fft_len=fft_len, mel_nbins=mel_nbins,
           mel_lower_hz=mel_lower_hz, mel_upper_hz=mel_upper_hz)
    return _frames(data, nSamples, frame_count, capi.MFCC_F32_MCU_LOG if use_log else capi.MFCC_F32_MCU)
