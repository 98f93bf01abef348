"""-m gpu: K1's latency instantiation (groups of 1, 2 or 4 frames for launches that would not fill the machine's warps with groups
of 8, mfcc_q15_kernel.cu launch_mfcc_q15) against the batch instantiation and the oracle, at every frame count where the group
size or the grid shape changes."""
import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu

SLOTS = 148 * 20      # warp slots of a B200: 1 frame per group up to here, 2 up to 2 * SLOTS, 4 up to 4 * SLOTS, then 8


def test_small_launches_equal_the_batch_path(edison, oracle):
    n_all = 4 * SLOTS + 700
    pcm = common.synth_frames(n_all, seed=4242)
    full = edison.mfcc_q15_batch(pcm)                      # one launch, groups of 8
    _, full_ni = edison.mfcc_q15_clips(pcm.reshape(-1), n_all, 1, 1024, want_mfcc=False)
    want = oracle.mfcc(pcm[:600])
    assert (full[:600] == want).all()
    assert (full_ni == np.clip(full[:, :13], -128, 127).astype(np.int8)).all()
    for n in (1, 2, 3, 4, 5, 31, 32, 33, 591, 592, 593, 1024, SLOTS - 1, SLOTS, SLOTS + 1, SLOTS + 2, 2 * SLOTS, 2 * SLOTS + 1,
              2 * SLOTS + 3, 4 * SLOTS - 1, 4 * SLOTS, 4 * SLOTS + 1):
        got = edison.mfcc_q15_batch(pcm[:n])
        assert (got == full[:n]).all(), n
        _, ni = edison.mfcc_q15_clips(pcm[:n].reshape(-1), n, 1, 1024, want_mfcc=False)
        assert (ni == full_ni[:n]).all(), n
    # odd clip stride: every second frame starts at an odd sample offset -> the unaligned instantiation of the small path
    flat = np.zeros(699 * 1025 + 1024, np.int16)
    for i in range(700):
        flat[i * 1025:i * 1025 + 1024] = pcm[i]
    mf, _ = edison.mfcc_q15_clips(flat, 700, 1, 1025, want_net_in=False)
    assert (mf.reshape(700, 32) == full[:700]).all()
