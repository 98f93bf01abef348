"""-m gpu: out-of-bounds evidence without compute-sanitizer (closed on this pool: profiles/r02_sanitizer_closed.txt).

Every kernel family runs on buffers carved out of larger device allocations.  Around every input and output lies a guard band of
canary bytes.  After a launch (a) the outputs are exact, (b) every guard band is untouched -- no out-of-bounds store --, and (c) the
outputs do not change when the guards around the inputs are re-poisoned with other bytes -- no out-of-bounds load that matters.
Sizes hit ragged tails: partial warp groups in K1, partial 32-clip batches and the unaligned byte-load path in K2."""
import ctypes

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
G = 4096   # guard bytes on each side


class Carved:
    def __init__(self, nbytes, dev, canary=0xC3, shift=0):
        self.nbytes, self.shift = nbytes, shift
        self.whole = torch.full((G + shift + nbytes + G,), canary, dtype=torch.uint8, device=dev)
        self.canary = canary

    @property
    def view(self):
        return self.whole[G + self.shift: G + self.shift + self.nbytes]

    @property
    def ptr(self):
        return ctypes.c_void_p(self.whole.data_ptr() + G + self.shift)

    def put(self, arr):
        self.view.copy_(torch.from_numpy(np.ascontiguousarray(arr).view(np.uint8).reshape(-1)))

    def get(self, dtype, shape):
        return self.view.cpu().numpy().view(dtype).reshape(shape).copy()

    def repoison(self, canary):
        self.whole[:G + self.shift] = canary
        self.whole[G + self.shift + self.nbytes:] = canary
        self.canary = canary

    def intact(self):
        a = self.whole[:G + self.shift]
        b = self.whole[G + self.shift + self.nbytes:]
        return bool((a == self.canary).all().item() and (b == self.canary).all().item())


@pytest.fixture(scope="module")
def dev(edison):
    return torch.device("cuda", torch.cuda.current_device())


def sync(edison):
    edison.capi.check(edison.capi.lib().edison_sync(None))


@pytest.mark.parametrize("n_frames", [1, 9, 333])
def test_k1_guards(edison, oracle, dev, n_frames):
    L, chk = edison.capi.lib(), edison.capi.check
    pcm = common.synth_frames(n_frames, seed=n_frames)
    want = oracle.mfcc(pcm)
    cin, cm, cn = Carved(pcm.nbytes, dev), Carved(n_frames * 64, dev), Carved(n_frames * 13, dev)
    cin.put(pcm)
    outs = []
    for canary in (0xC3, 0x5A):
        cin.repoison(canary)
        chk(L.edison_mfcc_q15_clips_dev(cin.ptr, n_frames, 1, 1024, cm.ptr, cn.ptr, None))
        sync(edison)
        assert cin.intact() and cm.intact() and cn.intact(), "guard band written"
        outs.append((cm.get(np.int16, (n_frames, 32)), cn.get(np.int8, (n_frames, 13))))
    assert (outs[0][0] == want).all() and (outs[0][1] == oracle.net_input(want)).all()
    assert (outs[0][0] == outs[1][0]).all() and (outs[0][1] == outs[1][1]).all(), "result depends on bytes outside the input"


@pytest.mark.parametrize("n,shift", [(1, 0), (31, 0), (33, 0), (100, 0), (33, 1), (64, 3), (97, 16)])
def test_k2_guards(edison, oracle, dev, n, shift):
    """shift != 0 (mod 16): the caller's buffer is not 16-byte aligned -> K2 loads its input with byte loads instead of bulk copies"""
    L, chk = edison.capi.lib(), edison.capi.check
    x = np.random.default_rng(n + shift).integers(-128, 128, (n, 403)).astype(np.int8)
    wl, ws = oracle.forward(x)
    cin, cl, cs, cc = Carved(x.nbytes, dev, shift=shift), Carved(n * 10, dev, shift=shift), Carved(n * 10, dev), Carved(40, dev)
    cin.put(x)
    outs = []
    for canary in (0xC3, 0x5A):
        cin.repoison(canary)
        cc.view.zero_()
        chk(L.edison_infer_q7_counts_dev(cin.ptr, n, cl.ptr, cs.ptr, cc.ptr, None))
        sync(edison)
        assert cin.intact() and cl.intact() and cs.intact() and cc.intact(), "guard band written"
        outs.append((cl.get(np.int8, (n, 10)), cs.get(np.int8, (n, 10)), cc.get(np.int32, (10,))))
    assert (outs[0][0] == wl).all() and (outs[0][1] == ws).all()
    assert (outs[0][2] == np.bincount(wl.argmax(axis=1), minlength=10)).all()
    assert all((a == b).all() for a, b in zip(outs[0], outs[1])), "result depends on bytes outside the input"


def test_float_kernels_guards(edison, dev):
    """K3 (fp32 / fp64 float MFCC) and K4 (float CNN): guard bands intact, results independent of the input guards"""
    L, chk = edison.capi.lib(), edison.capi.check
    nf = 77
    pcm = common.synth_frames(nf, seed=8)
    cin = Carved(pcm.nbytes, dev)
    cin.put(pcm)
    for fn, esz, dt in ((L.edison_mfcc_f32_clips_dev, 4, np.float32), (L.edison_mfcc_f64_clips_dev, 8, np.float64)):
        co = Carved(nf * 32 * esz, dev)
        res = []
        for canary in (0xC3, 0x5A):
            cin.repoison(canary)
            chk(fn(cin.ptr, nf, 1, 1024, co.ptr, 0, None))
            sync(edison)
            assert cin.intact() and co.intact()
            res.append(co.get(dt, (nf, 32)))
        assert np.array_equal(res[0], res[1]) and np.isfinite(res[0]).all()
    n = 37
    x = np.random.default_rng(1).normal(0, 30, (n, 403)).astype(np.float32)
    cx, cl, cs = Carved(x.nbytes, dev), Carved(n * 40, dev), Carved(n * 40, dev)
    cx.put(x)
    res = []
    for canary in (0xC3, 0x5A):
        cx.repoison(canary)
        chk(L.edison_infer_f32_batch_dev(cx.ptr, n, cl.ptr, cs.ptr, None))
        sync(edison)
        assert cx.intact() and cl.intact() and cs.intact()
        res.append((cl.get(np.float32, (n, 10)), cs.get(np.float32, (n, 10))))
    assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])


@pytest.mark.parametrize("incremental", [0, 1])
def test_stream_hop_guards(edison, dev, incremental):
    L, chk = edison.capi.lib(), edison.capi.check
    n = 37
    cin, cs, cf, cd = Carved(n * 2048, dev), Carved(n * 10, dev), Carved(n * 40, dev), Carved(n * 4, dev)
    with edison.Streams(n, incremental) as st:
        for h in range(40):
            cin.put(common.synth_frames(n, seed=300 + h))
            chk(L.edison_streams_push_hop_dev(st.handle, cin.ptr, cs.ptr, cf.ptr, cd.ptr, None))
            sync(edison)
            assert cin.intact() and cs.intact() and cf.intact() and cd.intact(), "hop %d wrote a guard band" % h
