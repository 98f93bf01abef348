"""-m gpu: K2 (tcgen05 pipeline: bulk copies, mbarrier slot ring, cross-batch unit order) over many launch shapes -- every clip count
from 1 to 70 (all ragged-tail cases of the 32-clip batches, one and several batches per CTA start), counts around multiples of 32 x 148
(the grid-filling boundaries), random counts up to 40,000, and input pointers at every byte alignment (the bulk-copy path needs 16 bytes;
the others take the guarded loads).  Expectations come from the oracle through a seeded gather of a pool, so each launch is checked on
all of its clips, logits and softmax, bit for bit."""
import ctypes

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def P(t):
    return ctypes.c_void_p(t.data_ptr())


def test_k2_launch_shape_sweep(edison, oracle):
    pool = 512
    rng = np.random.default_rng(77)
    base = oracle.net_input(oracle.mfcc(common.synth_frames(pool * 31, seed=78))).reshape(pool, 403)
    base[:64] = rng.integers(-128, 128, (64, 403)).astype(np.int8)        # arbitrary int8 inputs too, not only MFCC-shaped ones
    want_lg, want_sm = oracle.forward(base)
    dev = torch.device("cuda")
    L = edison.lib()
    tbase = torch.from_numpy(base).to(dev)
    sizes = list(range(1, 71)) + [32 * 148 + d for d in (-33, -32, -1, 0, 1, 31, 32, 33)] + [2 * 32 * 148 + d for d in (-1, 0, 1)]
    sizes += [int(v) for v in rng.integers(71, 40000, 60)]
    g = torch.Generator().manual_seed(5)
    for it, n in enumerate(sizes):
        idx = torch.randint(0, pool, (n,), generator=g)
        off = it % 16                                     # byte misalignment of the input pointer
        buf = torch.empty(n * 403 + 16, dtype=torch.int8, device=dev)
        x = buf[off:off + n * 403]
        x.copy_(tbase[idx.to(dev)].reshape(-1))
        lg = torch.full((n, 10), 99, dtype=torch.int8, device=dev)
        sm = torch.full((n, 10), 99, dtype=torch.int8, device=dev)
        torch.cuda.synchronize()
        edison.capi.check(L.edison_infer_q7_batch_dev(P(x), n, P(lg), P(sm), None))
        edison.capi.check(L.edison_sync(None))
        ix = idx.numpy()
        assert np.array_equal(lg.cpu().numpy(), want_lg[ix]), "logits, n = %d, offset %d" % (n, off)
        assert np.array_equal(sm.cpu().numpy(), want_sm[ix]), "softmax, n = %d, offset %d" % (n, off)
