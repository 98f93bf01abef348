"""-m gpu: parity at BASELINE.json's full single-GPU sizes, through a size-independent property: the inputs are built on the
device by gathering a small pool of oracle-checked clips in a seeded random order, so every one of the 131,072 (config #3
shard) / 983,040 (config #2) results has a known bit-exact expectation that cost only `pool` oracle evaluations.  Also
checked: batch-composition independence (one launch vs. ragged sub-launches) and run-to-run determinism."""
import ctypes

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def P(t):
    return ctypes.c_void_p(t.data_ptr())


def test_config3_shard_131072_clips_bit_exact(edison, oracle):
    n, pool, clip = 131072, 192, 31 * 1024
    base = common.synth_frames(pool * 31, seed=303).reshape(pool, clip)          # all six distributions, saturating ones included
    want_lg, want_sm, want_ni = oracle.kws(base, pool)
    dev = torch.device("cuda")
    g = torch.Generator(device="cpu").manual_seed(7)
    idx = torch.randint(0, pool, (n,), generator=g)
    pcm = torch.from_numpy(base).to(dev)[idx.to(dev)].contiguous()               # 8.3 GB on the device
    lg = torch.empty(n, 10, dtype=torch.int8, device=dev)
    sm = torch.empty(n, 10, dtype=torch.int8, device=dev)
    ni = torch.empty(n, 403, dtype=torch.int8, device=dev)
    L = edison.lib()
    torch.cuda.synchronize()      # the library launches on its own stream: torch's gather must have finished
    edison.capi.check(L.edison_kws_batch_dev(P(pcm), n, clip, P(lg), P(sm), P(ni), None))
    edison.capi.check(L.edison_sync(None))
    idx_np = idx.numpy()
    assert np.array_equal(lg.cpu().numpy(), want_lg[idx_np])
    assert np.array_equal(sm.cpu().numpy(), want_sm[idx_np])
    assert np.array_equal(ni.cpu().numpy(), want_ni[idx_np])
    # determinism + batch-composition independence: ragged sub-launches reproduce the single launch
    lg2 = torch.empty_like(lg)
    ni2 = torch.empty_like(ni)          # the _dev variant needs the [n][403] scratch between K1 and K2
    off = 0
    for m in (1, 15, 16, 17, 4099, 50000, n - 1 - 15 - 16 - 17 - 4099 - 50000):
        edison.capi.check(L.edison_kws_batch_dev(P(pcm[off:]), m, clip, P(lg2[off:]), None, P(ni2[off:]), None))
        off += m
    edison.capi.check(L.edison_sync(None))
    assert off == n and torch.equal(lg, lg2) and torch.equal(ni, ni2)


def test_config2_983040_frames_bit_exact(edison, oracle):
    n_clips, fpc, stride, pool = 65536, 15, 16000, 96
    rng = np.random.default_rng(11)
    base = np.zeros((pool, stride), dtype=np.int16)
    base[:, :15 * 1024] = common.synth_frames(pool * 15, seed=404).reshape(pool, 15 * 1024)
    base[:, 15 * 1024:] = rng.integers(-32768, 32767, (pool, stride - 15 * 1024))   # the 640 unused tail samples must not matter
    want = oracle.mfcc(base[:, :15 * 1024].reshape(-1, 1024)).reshape(pool, 15, 32)
    dev = torch.device("cuda")
    idx = torch.randint(0, pool, (n_clips,), generator=torch.Generator().manual_seed(9))
    pcm = torch.from_numpy(base).to(dev)[idx.to(dev)].contiguous()
    mf = torch.empty(n_clips * fpc, 32, dtype=torch.int16, device=dev)
    L = edison.lib()
    torch.cuda.synchronize()
    edison.capi.check(L.edison_mfcc_q15_clips_dev(P(pcm), n_clips, fpc, stride, P(mf), None, None))
    edison.capi.check(L.edison_sync(None))
    got = mf.cpu().numpy().reshape(n_clips, 15, 32)
    assert np.array_equal(got, want[idx.numpy()])
    # checksum of checksums: a second run and a run that also writes the net-input rows give the same MFCCs
    mf2 = torch.empty_like(mf)
    ni = torch.empty(n_clips * fpc, 13, dtype=torch.int8, device=dev)
    edison.capi.check(L.edison_mfcc_q15_clips_dev(P(pcm), n_clips, fpc, stride, P(mf2), P(ni), None))
    edison.capi.check(L.edison_sync(None))
    assert torch.equal(mf, mf2)
    assert torch.equal(ni.to(torch.int16), mf2[:, :13].clamp(-128, 127))            # mfccToNetInput, app.c:684-693


def test_config4_262144_float_clips_vs_restatement(edison, oracle):
    """float CNN at config #4's size: every one of the 262,144 results is compared with the fp64 restatement (oracle/cube_float.py) of
    the same input, within the tolerance of tests/test_cube_float.py (softmax max-abs 1e-4, logits 1e-3 relative) -- the inputs are a
    seeded gather of a pool whose restatement outputs cost 160 CPU evaluations; and the large launch equals a small-batch launch of
    the same inputs bit for bit (per-clip arithmetic order does not depend on the batch)"""
    import os
    import sys
    sys.path.insert(0, os.path.join(common.ROOT, "oracle"))
    import cube_float as cf
    n, pool = 262144, 160
    x = oracle.mfcc(common.synth_frames(pool * 31, seed=505))[:, :13].astype(np.float32).reshape(pool, 403)
    want_lg, want_sm = cf.forward(cf.load(os.path.join(common.ROOT, "edison_b200", "data", "kws_cube_f32.bin")), x)
    lg_small, sm_small = edison.infer_f32_batch(x)
    dev = torch.device("cuda")
    idx = torch.randint(0, pool, (n,), generator=torch.Generator().manual_seed(3))
    xin = torch.from_numpy(x).to(dev)[idx.to(dev)].contiguous()
    lg = torch.empty(n, 10, dtype=torch.float32, device=dev)
    sm = torch.empty(n, 10, dtype=torch.float32, device=dev)
    L = edison.lib()
    torch.cuda.synchronize()
    edison.capi.check(L.edison_infer_f32_batch_dev(P(xin), n, P(lg), P(sm), None))
    edison.capi.check(L.edison_sync(None))
    lg, sm, ix = lg.cpu().numpy(), sm.cpu().numpy(), idx.numpy()
    assert np.abs(sm - want_sm[ix]).max() <= 1e-4
    assert np.abs(lg - want_lg[ix]).max() <= 1e-3 * max(1.0, np.abs(want_lg).max())
    assert np.array_equal(lg, lg_small[ix])      # same kernel, same per-clip arithmetic order: identical bits
    assert np.array_equal(sm, sm_small[ix])
