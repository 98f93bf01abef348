"""NNoM models with general graphs (version-2 blobs; include/edison_b200.h section 2d).
Pin: upstream NNoM example models that the reference tree carries (mnist-simple: SAME convs and pools, two dense layers;
uci-inception: strided 1xk convs, parallel branches, Concat; keyword_spotting again through the graph converter), run through
the reference's UNMODIFIED NNoM runtime built for the host (tests/golden/nnom_graphs.npz, make_golden_nnom_graphs.py).
CPU: the oracle's graph forward against those fixtures.  GPU: edison_model_* against the fixtures and against the oracle on
random graphs that mix every supported node type."""
import ctypes
import os
import struct

import numpy as np
import pytest

import common

GOLD = os.path.join(common.ROOT, "tests", "golden")
MODELS = {"mnist_simple": ((28, 28, 1), 10), "uci_inception": ((1, 128, 9), 6), "keyword_spotting": ((63, 12, 1), 35)}
OP = dict(input=0, conv=1, dense=2, maxpool=3, concat=4, softmax=5, gmaxpool=6, add=7)


def blob_path(key):
    return os.path.join(GOLD, "nnom_graph_%s.bin" % key)


def oracle_graph(oracle, blob, x, mode):
    L = oracle.L
    L.eo_graph_node_output.restype = ctypes.c_void_p
    assert L.eo_graph_load(os.fsencode(blob)) == 0
    ni, no = L.eo_graph_in_size(), L.eo_graph_out_size()
    x = np.ascontiguousarray(x, dtype=np.int8).reshape(-1, ni)
    lg = np.zeros((x.shape[0], no), dtype=np.int8)
    sm = np.zeros((x.shape[0], no), dtype=np.int8)
    assert L.eo_graph_forward(common._p(x), ctypes.c_size_t(x.shape[0]), common._p(lg), common._p(sm), ctypes.c_int(mode)) == 0
    return lg, sm


@pytest.mark.parametrize("key", sorted(MODELS))
def test_oracle_graph_vs_reference_runtime(oracle, key):
    g = np.load(os.path.join(GOLD, "nnom_graphs.npz"))
    lg, sm = oracle_graph(oracle, blob_path(key), g[key + "_x"], 2)      # NNoM local backend -> local_softmax_q7
    assert np.array_equal(lg, g[key + "_logits"])
    assert np.array_equal(sm, g[key + "_softmax_local"])
    assert len(set(map(bytes, g[key + "_logits"]))) > g[key + "_x"].shape[0] // 2      # the fixture is not degenerate


def test_oracle_graph_equals_chain_oracle_on_the_chain_model(oracle):
    """the same header through both converters (version 1 and version 2 blobs) and both oracle paths"""
    from test_nnom_generic import EXAMPLE, oracle_generic
    g = np.load(os.path.join(GOLD, "nnom_example.npz"))
    for mode in (0, 1, 2):
        a = oracle_generic(oracle, EXAMPLE, g["x"], mode)
        b = oracle_graph(oracle, blob_path("keyword_spotting"), g["x"], mode)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_oracle_graph_rejects_bad_blobs(oracle, tmp_path):
    L = oracle.L
    raw = open(blob_path("uci_inception"), "rb").read()
    p = tmp_path / "t.bin"
    p.write_bytes(raw[:len(raw) // 2])
    assert L.eo_graph_load(os.fsencode(str(p))) == -5
    p.write_bytes(raw[:8] + struct.pack("<I", 3) + raw[12:])
    assert L.eo_graph_load(os.fsencode(str(p))) == -4
    assert L.eo_graph_load(os.fsencode(blob_path("uci_inception"))) == 0


class GraphWriter:
    """writes version-2 blobs node by node, tracking shapes the way the converter does"""

    def __init__(self, rng, in_shape):
        self.rng, self.nodes, self.shapes = rng, [], []
        self._add("input", [], list(in_shape), tuple(in_shape))

    def _add(self, op, ins, p, shape, w=None, b=None):
        self.nodes.append((op, ins, p, w, b))
        self.shapes.append(shape)
        return len(self.nodes) - 1

    @staticmethod
    def _dim(n, k, s, same):
        return -(-n // s) if same else -(-(n - k + 1) // s)

    def conv(self, src, cout, kh, kw, sh=1, sw=1, same=0, relu=1):
        h, w, c = self.shapes[src]
        wt = self.rng.integers(-40, 40, cout * kh * kw * c).astype(np.int8)
        b = self.rng.integers(-60, 60, cout).astype(np.int8)
        p = [cout, kh, kw, sh, sw, same, int(self.rng.integers(5, 9)), int(self.rng.integers(0, 5)), relu, c]
        return self._add("conv", [src], p, (self._dim(h, kh, sh, same), self._dim(w, kw, sw, same), cout), wt, b)

    def pool(self, src, kh, kw, sh, sw, same=0):
        h, w, c = self.shapes[src]
        return self._add("maxpool", [src], [kh, kw, sh, sw, same], (self._dim(h, kh, sh, same), self._dim(w, kw, sw, same), c))

    def gpool(self, src):
        return self._add("gmaxpool", [src], [], (1, 1, self.shapes[src][2]))

    def concat(self, srcs):
        h, w, c = self.shapes[srcs[0]]
        return self._add("concat", list(srcs), [], (h, w, c * len(srcs)))

    def add(self, a, b, rshift=1, relu=0):
        return self._add("add", [a, b], [rshift, relu], self.shapes[a])

    def dense(self, src, cout, relu=0):
        h, w, c = self.shapes[src]
        n_in = h * w * c
        wt = self.rng.integers(-30, 30, cout * n_in).astype(np.int8)
        b = self.rng.integers(-60, 60, cout).astype(np.int8)
        return self._add("dense", [src], [cout, n_in, int(self.rng.integers(6, 10)), int(self.rng.integers(0, 4)), relu], (1, 1, cout), wt, b)

    def softmax(self, src):
        return self._add("softmax", [src], [], self.shapes[src])

    def write(self, path):
        out = bytearray(b"EDNNQ7\0\0" + struct.pack("<II", 2, len(self.nodes)))
        for op, ins, p, w, b in self.nodes:
            out += struct.pack("<II8I10I", OP[op], len(ins), *(ins + [0xFFFFFFFF] * (8 - len(ins))), *(p + [0] * (10 - len(p))))
            if w is not None:
                out += w.tobytes() + b.tobytes()
                out += b"\0" * (-len(out) % 4)
        out += struct.pack("<iii", 1, -128, 127)
        open(path, "wb").write(out)


def random_graph(case, rng, path):
    if case == 0:      # residual block with SAME convs, odd sizes, a pool that pads, global max pool
        g = GraphWriter(rng, (15, 11, 3))
        a = g.conv(0, 8, 3, 3, same=1)
        b = g.conv(a, 8, 3, 3, same=1, relu=0)
        s = g.add(a, b, rshift=1, relu=1)
        p = g.pool(s, 3, 3, 2, 2, same=1)
        c = g.conv(p, 12, 5, 1, same=1)
        d = g.dense(g.gpool(c), 9)
        g.softmax(d)
    elif case == 1:    # inception-like: strided stem, three branches of different kernel widths, concat, 4x4 pool with padding
        g = GraphWriter(rng, (24, 20, 4))
        st = g.conv(0, 16, 3, 5, sh=2, sw=1, same=1)
        b1 = g.conv(st, 8, 1, 1)
        b2 = g.conv(st, 8, 3, 3, same=1)
        b3 = g.pool(g.conv(st, 8, 7, 7, same=1, relu=0), 3, 3, 1, 1, same=1)
        c = g.concat([b2, b1, b3])
        p = g.pool(c, 4, 4, 4, 4, same=1)
        d = g.dense(p, 32, relu=1)
        g.softmax(g.dense(d, 5))
    elif case == 2:    # valid convs with strides in both axes, channels not a multiple of 4, no softmax node
        g = GraphWriter(rng, (31, 13, 1))
        a = g.conv(0, 6, 5, 5, sh=2, sw=1)
        b = g.pool(a, 2, 1, 2, 1)
        c = g.conv(b, 10, 3, 3, sh=1, sw=2, relu=0)
        g.dense(c, 11)
    else:              # a node read by several consumers, add of two pooled branches, 1-D shapes
        g = GraphWriter(rng, (1, 64, 8))
        a = g.conv(0, 16, 1, 9, sw=2, same=1)
        p1 = g.pool(a, 1, 2, 1, 2)
        p2 = g.pool(g.conv(a, 16, 1, 3, same=1), 1, 2, 1, 2)
        s = g.add(p1, p2, rshift=2)
        c = g.concat([s, p1])
        g.softmax(g.dense(c, 6))
    g.write(path)
    return g.shapes[0]


@pytest.mark.parametrize("case", range(4))
def test_random_graph_shapes_agree_with_oracle(oracle, tmp_path, case):
    """the writer's shape rules == the oracle loader's (a wrong rule would make dense sizes disagree and the load fail)"""
    blob = str(tmp_path / "g.bin")
    in_shape = random_graph(case, np.random.default_rng(300 + case), blob)
    assert oracle.L.eo_graph_load(os.fsencode(blob)) == 0
    assert oracle.L.eo_graph_in_size() == int(np.prod(in_shape))


@pytest.mark.gpu
@pytest.mark.parametrize("key", sorted(MODELS))
def test_gpu_graph_vs_reference_runtime(edison, key):
    g = np.load(os.path.join(GOLD, "nnom_graphs.npz"))
    edison.set_softmax_mode(edison.SOFTMAX_NNOM_LOCAL)
    try:
        with edison.Model(blob_path(key)) as m:
            assert (m.in_shape, m.n_out) == MODELS[key]
            lg, sm = m.infer(g[key + "_x"])
            m.set_naive(True)
            lg_n, sm_n = m.infer(g[key + "_x"])
            assert np.array_equal(lg, lg_n) and np.array_equal(sm, sm_n)
    finally:
        edison.set_softmax_mode(edison.SOFTMAX_CMSIS_SIMD)
    assert np.array_equal(lg, g[key + "_logits"])
    assert np.array_equal(sm, g[key + "_softmax_local"])


@pytest.mark.gpu
@pytest.mark.parametrize("case", range(4))
def test_gpu_random_graphs_vs_oracle(edison, oracle, tmp_path, case):
    rng = np.random.default_rng(300 + case)
    blob = str(tmp_path / "g.bin")
    in_shape = random_graph(case, rng, blob)
    n = 29
    x = rng.integers(-128, 128, (n, int(np.prod(in_shape)))).astype(np.int8)
    x[0] = -128
    for mode in (0, 1, 2):
        want_lg, want_sm = oracle_graph(oracle, blob, x, mode)
        edison.set_softmax_mode(mode)
        try:
            with edison.Model(blob) as m:
                lg, sm = m.infer(x)
                lg1, sm1 = m.infer(x[:1])                # a different batch size re-bases every arena offset
                m.set_naive(True)                        # every node on the one-thread-per-output kernels: same bits
                lg_n, sm_n = m.infer(x)
        finally:
            edison.set_softmax_mode(edison.SOFTMAX_CMSIS_SIMD)
        assert np.array_equal(lg, want_lg), "logits, mode %d" % mode
        assert np.array_equal(sm, want_sm), "softmax, mode %d" % mode
        assert np.array_equal(lg1, want_lg[:1]) and np.array_equal(sm1, want_sm[:1])
        assert np.array_equal(lg_n, want_lg) and np.array_equal(sm_n, want_sm)
    assert np.unique(want_lg).size > 8


@pytest.mark.gpu
def test_gpu_graph_rejects_bad_graphs(edison, tmp_path):
    L = edison.lib()
    h = ctypes.c_void_p()
    rng = np.random.default_rng(1)
    g = GraphWriter(rng, (8, 8, 4))
    a = g.conv(0, 8, 3, 3, same=1)
    b = g.conv(0, 4, 3, 3, same=1)
    g.concat([a, b])                                       # unequal channel counts: not reproducible in the reference
    g.dense(3, 5)
    p = str(tmp_path / "bad.bin")
    g.write(p)
    assert L.edison_model_load(os.fsencode(p), ctypes.byref(h)) == -6
    assert b"concat" in L.edison_last_error()
    raw = open(blob_path("mnist_simple"), "rb").read()
    open(p, "wb").write(raw[:5000])
    assert L.edison_model_load(os.fsencode(p), ctypes.byref(h)) == -4
