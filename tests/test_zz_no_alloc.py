"""-m gpu: edison_kws_batch_noalloc_dev (include/edison_b200.h; no counterpart in the reference):
K1 -> K2 with the [n][403] scratch in a static device array instead of a caller buffer / cudaMalloc.

Checked through the C ABI against edison_kws_batch_dev on the same inputs, bit for bit (the tolerance of every integer test
here), on two calls in a row of different sizes and contents -- the second one smaller and ragged, so stale rows of the first
call sit right behind its data in the shared array.  Collected last on purpose
(the name sorts after every other test file)."""
import ctypes

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

MAX_CLIPS = 8192          # EDISON_NOALLOC_MAX_CLIPS
CLIP = 31 * 1024


def P(t):
    return ctypes.c_void_p(t.data_ptr())


def _clips(n, pool, seed, dev):
    base = common.synth_frames(pool * 31, seed=seed).reshape(pool, CLIP)
    idx = torch.randint(0, pool, (n,), generator=torch.Generator(device="cpu").manual_seed(seed))
    return torch.from_numpy(base).to(dev)[idx.to(dev)].contiguous()


def _existing(edison, pcm, n):
    L = edison.lib()
    lg = torch.empty(n, 10, dtype=torch.int8, device=pcm.device)
    sm = torch.empty(n, 10, dtype=torch.int8, device=pcm.device)
    ni = torch.empty(n, 403, dtype=torch.int8, device=pcm.device)
    edison.capi.check(L.edison_kws_batch_dev(P(pcm), n, CLIP, P(lg), P(sm), P(ni), None))
    edison.capi.check(L.edison_sync(None))
    return lg, sm


def test_noalloc_matches_existing_entry_point_twice(edison):
    dev = torch.device("cuda")
    L = edison.lib()
    assert hasattr(L, "edison_kws_batch_noalloc_dev")
    sizes = (MAX_CLIPS, 333)                      # the maximum, then a ragged batch (333 = 10 * 32 + 13)
    pcm = [_clips(n, 48, 900 + i, dev) for i, n in enumerate(sizes)]
    want = [_existing(edison, p, n) for p, n in zip(pcm, sizes)]
    got = [(torch.full((n, 10), 99, dtype=torch.int8, device=dev), torch.full((n, 10), 99, dtype=torch.int8, device=dev)) for n in sizes]
    # set-up call: the first use on a device looks the array's address up, which makes the driver load the module that holds the
    # 3.15 MiB array (lazy module loading); from here on the free device memory must not move
    warm = torch.empty(1, 10, dtype=torch.int8, device=dev)
    edison.capi.check(L.edison_kws_batch_noalloc_dev(P(pcm[1]), 1, CLIP, P(warm), None, None))
    edison.capi.check(L.edison_sync(None))
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for p, n, (lg, sm) in zip(pcm, sizes, got):   # two calls in a row on the library's stream, no host sync in between
        edison.capi.check(L.edison_kws_batch_noalloc_dev(P(p), n, CLIP, P(lg), P(sm), None))
    edison.capi.check(L.edison_sync(None))
    free1 = torch.cuda.mem_get_info()[0]
    for (wl, ws), (lg, sm), n in zip(want, got, sizes):
        assert torch.equal(lg, wl), "logits differ at n = %d" % n
        assert torch.equal(sm, ws), "softmax differs at n = %d" % n
    assert free1 == free0, "device memory in use changed across the calls: %d -> %d bytes free" % (free0, free1)
    # once more in the opposite order, logits only, on a caller's stream
    st = torch.cuda.Stream()
    lg_small = torch.full((sizes[1], 10), 99, dtype=torch.int8, device=dev)
    lg_big = torch.full((sizes[0], 10), 99, dtype=torch.int8, device=dev)
    torch.cuda.synchronize()
    edison.capi.check(L.edison_kws_batch_noalloc_dev(P(pcm[1]), sizes[1], CLIP, P(lg_small), None, ctypes.c_void_p(st.cuda_stream)))
    edison.capi.check(L.edison_kws_batch_noalloc_dev(P(pcm[0]), sizes[0], CLIP, P(lg_big), None, ctypes.c_void_p(st.cuda_stream)))
    st.synchronize()
    assert torch.equal(lg_small, want[1][0]) and torch.equal(lg_big, want[0][0])


def test_noalloc_rejects_oversize_and_bad_arguments(edison):
    dev = torch.device("cuda")
    L = edison.lib()
    L.edison_last_error.restype = ctypes.c_char_p
    pcm = torch.zeros(CLIP, dtype=torch.int16, device=dev)
    out = torch.full((1, 10), 99, dtype=torch.int8, device=dev)
    E_ARG = -1
    # beyond the static scratch: refused before anything is launched or read
    assert L.edison_kws_batch_noalloc_dev(P(pcm), MAX_CLIPS + 1, CLIP, P(out), None, None) == E_ARG
    msg = L.edison_last_error().decode()
    assert "8192" in msg and "8193" in msg, msg
    assert L.edison_kws_batch_noalloc_dev(None, 1, CLIP, P(out), None, None) == E_ARG
    assert L.edison_kws_batch_noalloc_dev(P(pcm), 1, CLIP, None, None, None) == E_ARG
    assert L.edison_kws_batch_noalloc_dev(P(pcm), 2, 0, P(out), None, None) == E_ARG
    assert L.edison_kws_batch_noalloc_dev(P(pcm), 0, CLIP, None, None, None) == 0      # nothing to do
    edison.capi.check(L.edison_sync(None))
    assert (out == 99).all()


def test_noalloc_call_can_be_captured_into_a_cuda_graph(edison):
    """what the entry point is for: under stream capture (where cudaMalloc is illegal) the call records its two kernels; replays on new
    inputs give the existing entry point's bytes"""
    dev = torch.device("cuda")
    L = edison.lib()
    n = 1000
    pcm = _clips(n, 48, 931, dev)
    lg = torch.full((n, 10), 99, dtype=torch.int8, device=dev)
    sm = torch.full((n, 10), 99, dtype=torch.int8, device=dev)
    edison.capi.check(L.edison_kws_batch_noalloc_dev(P(pcm), 1, CLIP, P(lg), None, None))     # set-up call before the capture
    edison.capi.check(L.edison_sync(None))
    graph, side = torch.cuda.CUDAGraph(), torch.cuda.Stream()
    with torch.cuda.graph(graph, stream=side):
        edison.capi.check(L.edison_kws_batch_noalloc_dev(P(pcm), n, CLIP, P(lg), P(sm), ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    for seed in (931, 932):                       # the first replay on the captured inputs, the second on new contents of the same buffer
        pcm.copy_(_clips(n, 48, seed, dev))
        want_lg, want_sm = _existing(edison, pcm, n)
        lg.fill_(99); sm.fill_(99)
        torch.cuda.synchronize()
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(lg, want_lg) and torch.equal(sm, want_sm), seed
