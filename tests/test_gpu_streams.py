"""Streaming scheduler (include/edison_b200.h section 3b) against a simulation of the firmware loop
(appAudioEvent app.c:635-661 + appMicMfccInfereContinuous app.c:288-394) built from the oracle's pieces."""
import numpy as np
import pytest

from common import synth_frames

pytestmark = pytest.mark.gpu


def simulate(oracle, pcm_hops, mode=1):
    """pcm_hops: int16 [hops, n, 1024].  Returns per hop (softmax, filt, detect) as the firmware would compute them."""
    hops, n, _ = pcm_hops.shape
    window = np.zeros((n, 31, 13), dtype=np.int8)
    filt = np.zeros((n, 10), dtype=np.float32)
    sm = np.zeros((n, 10), dtype=np.int8)
    out = []
    for h in range(hops):
        rows = oracle.net_input(oracle.mfcc(pcm_hops[h]))          # audioCalcMFCCs + clip to int8
        window[:, :-1] = window[:, 1:].copy()                      # mfccToNetInputPush
        window[:, -1] = rows
        if h + 1 > 31:
            _, sm = oracle.forward(window.reshape(n, 403), mode)
            # double arithmetic with the double literals 0.9 and (1.0 - 0.9), stored to float (app.c:35-39, 342)
            filt = (0.9 * filt.astype(np.float64) + (1.0 - 0.9) * sm.astype(np.float32).astype(np.float64)).astype(np.float32)
            idx = np.argmax(filt, axis=1)                          # arm_max_f32: first maximum, like np.argmax
            det = np.where(filt[np.arange(n), idx] > np.float32(0.5), idx, -1).astype(np.int32)
        else:
            det = np.full(n, -2, dtype=np.int32)
        out.append((sm.copy(), filt.copy(), det, window.copy()))
    return out


@pytest.mark.parametrize("incremental", [1, 0])
@pytest.mark.parametrize("n", [1, 37])
def test_streams_vs_firmware_loop(edison, oracle, n, incremental):
    hops = 58          # > 31 + 12 + 12: every ring of the incremental net wraps at least twice after inference starts
    pcm = synth_frames(hops * n, seed=77).reshape(hops, n, 1024)
    want = simulate(oracle, pcm)
    with edison.Streams(n, incremental) as st:
        for h in range(hops):
            sm, fl, dt = st.push_hop(pcm[h])
            w_sm, w_fl, w_dt, w_win = want[h]
            assert np.array_equal(dt, w_dt), "detect, hop %d" % h
            if h + 1 > 31:
                assert np.array_equal(sm, w_sm), "softmax, hop %d" % h
            assert np.array_equal(fl.view(np.uint32), w_fl.view(np.uint32)), "filter bits, hop %d" % h
            if h in (0, 12, 30, 31, 39, 57):
                for i in (0, n - 1):
                    assert np.array_equal(st.window(i), w_win[i]), "window, hop %d stream %d" % (h, i)
        assert st.hops == hops


def test_incremental_modes_equal_the_full_window_over_many_batches(edison):
    """5,017 streams = 157 batches of 32 on 148 CTAs (some CTAs carry two batches, the last batch is ragged): the incremental
    mode must give the full-window path's bytes at every hop (the full-window path is checked against the firmware loop above)"""
    n, hops = 5017, 52
    rng = np.random.default_rng(9)
    base = synth_frames(64 * 8, seed=9).reshape(8, 64, 1024)
    with edison.Streams(n, 0) as full, edison.Streams(n, 1) as inc:
        for h in range(hops):
            pcm = base[h % 8][rng.integers(0, 64, n)]
            a, b = full.push_hop(pcm), inc.push_hop(pcm)
            assert np.array_equal(a[2], b[2]), "detect, hop %d" % h
            if h + 1 > 31:
                assert np.array_equal(a[0], b[0]), "softmax, hop %d" % h
            assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), "filter bits, hop %d" % h


def test_incremental_mode_over_many_cache_rotations(edison):
    """the row cache rotates with period 12 hops (four phase sets x three slots): 150 hops on a ragged stream count, every hop compared
    with the full-window mode, through the pipelined submit / collect calls on one side and push_hop on the other"""
    n, hops = 97, 150
    rng = np.random.default_rng(21)
    base = synth_frames(48 * 6, seed=21).reshape(6, 48, 1024)
    with edison.Streams(n, 0) as full, edison.Streams(n, 1) as inc:
        pending, alive = None, []
        for h in range(hops):
            pcm = np.ascontiguousarray(base[h % 6][rng.integers(0, 48, n)])
            want = full.push_hop(pcm)
            alive = alive[-1:] + [inc.submit_hop(pcm)]      # a submitted hop's samples stay alive until it is collected
            if pending is not None:
                got = inc.collect_hop()
                for k in (0, 1, 2):
                    a, b = pending[k], got[k]
                    if k == 0 and h <= 31:
                        continue
                    assert np.array_equal(a.view(np.uint32) if a.dtype == np.float32 else a, b.view(np.uint32) if b.dtype == np.float32 else b), (h - 1, k)
            pending = want
        got = inc.collect_hop()
        assert np.array_equal(pending[2], got[2]) and np.array_equal(pending[0], got[0])


def test_streams_detects_when_softmax_saturates(edison, oracle):
    """Feed the same loud frame for many hops: the filter converges to the softmax output, so some class crosses 0.5
    exactly when the simulation says so (exercises the detect >= 0 branch)."""
    n, hops = 4, 80
    one = synth_frames(n, seed=5)
    pcm = np.broadcast_to(one, (hops, n, 1024)).copy()
    want = simulate(oracle, pcm)
    seen = False
    with edison.Streams(n) as st:
        for h in range(hops):
            _, fl, dt = st.push_hop(pcm[h])
            assert np.array_equal(dt, want[h][2])
            assert np.array_equal(fl.view(np.uint32), want[h][1].view(np.uint32))
            seen = seen or bool((dt >= 0).any())
    assert seen, "test input never crossed the threshold; pick another seed"


def test_streams_bad_arguments(edison):
    import ctypes
    L = edison.lib()
    h = ctypes.c_void_p()
    assert L.edison_streams_open(0, ctypes.byref(h)) == -1
    assert L.edison_streams_push_hop(None, None, None, None, None) == -1


def test_streams_pipelined_submit_collect(edison, oracle):
    """two hops in flight (copy of hop h + 1 overlaps the kernels of hop h): same results, in order, as the synchronous calls"""
    n, hops = 19, 45
    pcm = synth_frames(hops * n, seed=123).reshape(hops, n, 1024)
    want = simulate(oracle, pcm)
    got = []
    with edison.Streams(n) as st:
        keep = [st.submit_hop(pcm[0])]
        for h in range(1, hops):
            keep.append(st.submit_hop(pcm[h]))
            got.append(st.collect_hop())
            keep.pop(0)
        got.append(st.collect_hop())
        L = edison.lib()
        assert L.edison_streams_collect_hop(st.handle, None, None, None) == -1          # nothing in flight
        st.submit_hop(pcm[0]); st.submit_hop(pcm[1])
        assert L.edison_streams_submit_hop(st.handle, pcm[2].ctypes.data) == -1           # a third hop must wait for a collect
        st.collect_hop(); st.collect_hop()
    for h in range(hops):
        sm, fl, dt = got[h]
        assert np.array_equal(dt, want[h][2]), "detect, hop %d" % h
        assert np.array_equal(fl.view(np.uint32), want[h][1].view(np.uint32)), "filter bits, hop %d" % h
        if h + 1 > 31:
            assert np.array_equal(sm, want[h][0]), "softmax, hop %d" % h
