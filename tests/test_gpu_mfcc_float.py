"""K3 (float MFCC) against the reference's own outputs (tests/golden/mfcc_float.npz) and the numpy restatement.
Stated tolerance (north star / SURVEY 8(c)): max-abs <= 1e-3 and rmse <= 1e-4 on the coefficients.
  fp64 kernel: holds on every input, degenerate frames included; ~1e-9 on ordinary audio (asserted at 1e-7).
  fp32 kernel: holds on frames whose mel-band energies lie within 1e4 of each other (recorded audio, noise, speech-like
  synthetic); its rounding error scales with the frame's total energy, so a full-scale pure tone drowns the empty bands."""
import os
import sys

import numpy as np
import pytest

from common import synth_frames

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import mfcc_float as mf  # noqa: E402

pytestmark = pytest.mark.gpu
MAX_ABS, RMSE = 1e-3, 1e-4


def close(got, want, scale=1.0, max_abs=MAX_ABS, rmse=RMSE):
    d = got.astype(np.float64) - want
    assert np.abs(d).max(initial=0) <= max_abs * scale, "max-abs %g" % np.abs(d).max()
    assert np.sqrt(np.mean(d * d)) <= rmse * scale, "rmse %g" % np.sqrt(np.mean(d * d))


def ordinary(mel):
    """frames whose mel spectrum is nowhere (near) empty: log(mel + 1e-6) is then insensitive to rounding noise"""
    return mel.min(axis=1) > 1.0


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(ROOT, "tests", "golden", "mfcc_float.npz"))


@pytest.mark.parametrize("src", ["wav", "syn"])
def test_f64_host_mfcc_vs_reference(edison, gold, src):
    got = edison.mfcc_float_batch(gold[src + "_pcm"], edison.MFCC_F32_HOST)
    want = gold[src + "_host_mfcc"]
    assert got.shape == want.shape and got.dtype == np.float64
    close(got, want)                                        # every frame, degenerate ones included
    ok = ordinary(gold[src + "_host_mel"])
    assert ok.sum() >= 10
    close(got[ok], want[ok], max_abs=1e-7, rmse=1e-8)       # ordinary frames: numpy-equivalent


@pytest.mark.parametrize("src", ["wav", "syn"])
def test_f64_mcu_mfcc_vs_reference(edison, gold, src):
    want = gold[src + "_mcu_mfcc"]
    close(edison.mfcc_float_batch(gold[src + "_pcm"], edison.MFCC_F32_MCU), want, max_abs=1e-8, rmse=1e-9)
    close(edison.mfcc_float_batch(gold[src + "_pcm"], edison.MFCC_F32_MCU_LOG), gold[src + "_mculog_mfcc"])


@pytest.mark.parametrize("src", ["wav", "syn"])
def test_f32_vs_reference_on_ordinary_frames(edison, gold, src):
    pcm = gold[src + "_pcm"]
    mel = gold[src + "_host_mel"]
    ok = ordinary(mel) & (mel.min(axis=1) > 1e-4 * mel.max(axis=1))     # fp32: band energies within 1e4 of each other
    assert ok.sum() >= 10
    got = edison.mfcc_float_batch(pcm, edison.MFCC_F32_HOST, np.float32)
    assert got.dtype == np.float32
    close(got[ok], gold[src + "_host_mfcc"][ok])
    want = gold[src + "_mcu_mfcc"]
    # mfcc_mcu without log is linear in the spectrum (coefficients up to ~3e3): gate relative to that range
    close(edison.mfcc_float_batch(pcm, edison.MFCC_F32_MCU, np.float32), want, scale=max(1.0, np.abs(want).max()))


def test_reference_signature_mirror(edison, gold):
    from edison_b200 import mfcc_utils as mu
    pcm = gold["wav_pcm"]
    o = mu.mfcc(pcm, 16000, pcm.shape[0], 1024, 1024, 0, 1024, 32, 80.0, 7600.0)
    assert len(o) == 10 and o[3]["t_start"] == 3 * 1024 / 16000
    close(np.array([f["mfcc"] for f in o]), gold["wav_host_mfcc"], max_abs=1e-7, rmse=1e-8)
    o = mu.mfcc_mcu(pcm, 16000, pcm.shape[0], 1024, 1024, 0, 1024, 32, 80.0, 7600.0, 128)
    close(np.array([f["mfcc"] for f in o]), gold["wav_mcu_mfcc"], max_abs=1e-8, rmse=1e-9)
    with pytest.raises(ValueError):
        mu.mfcc(pcm, 8000, pcm.shape[0], 1024, 1024, 0, 1024, 32, 80.0, 7600.0)


@pytest.mark.parametrize("n", [1, 11, 12, 13, 4097])
def test_f64_vs_restatement_ragged(edison, n):
    pcm = synth_frames(n, seed=100 + n).reshape(-1)
    close(edison.mfcc_float_batch(pcm), mf.mfcc_host(pcm))


def test_clips_with_stride_and_odd_offset(edison):
    # config #2 framing: 1 s clips -> 15 frames, stride 16,000; then an odd sample stride (unaligned instantiation)
    n = 9
    pcm = synth_frames(n * 16 + 1, seed=9).reshape(-1)
    want = np.concatenate([mf.mfcc_host(pcm[c * 16000:c * 16000 + 15 * 1024]) for c in range(n)])
    close(edison.mfcc_float_clips(pcm, n, 15, 16000), want)
    want = np.concatenate([mf.mfcc_host(pcm[1 + c * 16001:1 + c * 16001 + 15 * 1024]) for c in range(n)])
    close(edison.mfcc_float_clips(pcm[1:], n, 15, 16001), want)
    close(edison.mfcc_float_clips(pcm[1:], n, 15, 16001, dtype=np.float32)[ordinary_rows(want)], want[ordinary_rows(want)])


def ordinary_rows(c):
    # coefficient 0 = sum of log-mels / 8: frames with empty bands sit far below the rest
    return c[:, 0] > 20.0


def test_empty(edison):
    assert edison.mfcc_float_batch(np.zeros(100, dtype=np.int16)).shape == (0, 32)
