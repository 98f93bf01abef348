"""-m gpu: the C ABI's edges -- lifetime, chunked host pipeline, pointer kinds, argument errors, independent stream sets."""
import ctypes

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu


def test_shutdown_and_reinit_give_the_same_bytes(edison, oracle):
    pcm = common.synth_frames(31 * 5, seed=1).reshape(5, -1)
    a = edison.kws_batch(pcm, 5)
    edison.shutdown()
    edison.init()
    b = edison.kws_batch(pcm, 5)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert np.array_equal(a[0], oracle.kws(pcm, 5)[0])


def test_small_chunks_exercise_the_two_slot_pipeline(edison, oracle):
    n = 53                                       # 7 full chunks of 7 and a ragged one of 4
    pcm = common.synth_frames(31 * n, seed=2).reshape(n, -1)
    edison.shutdown()
    edison.init(chunk_clips=7)
    try:
        lg, sm, ni = edison.kws_batch(pcm, n, want_net_in=True)
        mf, _ = edison.mfcc_q15_clips(pcm, n, 31, 31 * 1024, want_net_in=False)
    finally:
        edison.shutdown()
        edison.init()
    w_lg, w_sm, w_ni = oracle.kws(pcm, n)
    assert np.array_equal(lg, w_lg) and np.array_equal(sm, w_sm) and np.array_equal(ni, w_ni)
    assert np.array_equal(mf, oracle.mfcc(pcm.reshape(-1, 1024)))


def test_pageable_pinned_and_device_pointers_agree(edison):
    torch = pytest.importorskip("torch")
    L = edison.lib()
    frames = common.synth_frames(40, seed=3)
    want = edison.mfcc_q15_batch(frames)                                   # pageable numpy memory
    n_bytes = frames.nbytes
    L.edison_host_alloc.restype = ctypes.c_void_p
    p = L.edison_host_alloc(n_bytes)                                       # pinned
    assert p
    ctypes.memmove(p, frames.ctypes.data, n_bytes)
    out = np.zeros((40, 32), dtype=np.int16)
    edison.capi.check(L.edison_mfcc_q15_batch(ctypes.c_void_p(p), 40, out.ctypes.data_as(ctypes.c_void_p)))
    L.edison_host_free(ctypes.c_void_p(p))
    assert np.array_equal(out, want)
    d_in = torch.from_numpy(frames).cuda()                                 # device pointers, detected per call
    d_out = torch.zeros(40, 32, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    edison.capi.check(L.edison_mfcc_q15_batch(ctypes.c_void_p(d_in.data_ptr()), 40, ctypes.c_void_p(d_out.data_ptr())))
    assert np.array_equal(d_out.cpu().numpy(), want)


def test_argument_errors(edison):
    L = edison.lib()
    out = np.zeros((4, 10), dtype=np.int8)
    assert L.edison_kws_batch(None, 4, 31744, out.ctypes.data_as(ctypes.c_void_p), None, None) == -1
    assert L.edison_kws_batch(None, 0, 31744, None, None, None) == 0                 # nothing to do is not an error
    assert L.edison_set_softmax_mode(7) == -1
    assert b"softmax" in L.edison_last_error()
    assert L.edison_mfcc_q15_clips(np.zeros(8, np.int16).ctypes.data_as(ctypes.c_void_p), 1, 1, 0, None, None) in (0, -1)
    assert L.edison_infer_q7_batch(None, 3, out.ctypes.data_as(ctypes.c_void_p), None) == -1
    assert L.edison_launch_count() > 0


def test_two_stream_sets_are_independent(edison, oracle):
    hops, n = 36, 3
    a = common.synth_frames(hops * n, seed=4).reshape(hops, n, 1024)
    b = common.synth_frames(hops * n, seed=5).reshape(hops, n, 1024)
    with edison.Streams(n) as sa, edison.Streams(n) as sb, edison.Streams(n) as sa2:
        for h in range(hops):
            ra = sa.push_hop(a[h])
            sb.push_hop(b[h])
            ra2 = sa2.push_hop(a[h])
            assert all(np.array_equal(x, y) for x, y in zip(ra, ra2))
        assert not np.array_equal(sa.window(0), sb.window(0))
        assert np.array_equal(sa.window(1), sa2.window(1))
