"""ctypes binding of libedison_b200.so (include/edison_b200.h).  numpy arrays in, numpy arrays out."""
import ctypes
import os

import numpy as np

FRAME_LEN, N_MEL, N_COEFF, N_FRAMES, N_CLASSES = 1024, 32, 13, 31, 10
NET_IN = N_FRAMES * N_COEFF
SOFTMAX_CMSIS_REF, SOFTMAX_CMSIS_SIMD, SOFTMAX_NNOM_LOCAL = 0, 1, 2
MFCC_F32_HOST, MFCC_F32_MCU, MFCC_F32_MCU_LOG = 0, 1, 2

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


class EdisonError(RuntimeError):
    pass


class _Cfg(ctypes.Structure):
    _fields_ = [("device", ctypes.c_int), ("data_dir", ctypes.c_char_p), ("softmax_mode", ctypes.c_int),
                ("chunk_clips", ctypes.c_size_t)]


def lib_path():
    return os.path.join(_HERE, "libedison_b200.so")


def lib():
    """Load the shared library (no CUDA call is made by loading it)."""
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise EdisonError("%s is missing: run `make` (or __graft_entry__.build()) first; there is no fallback" % p)
        L = ctypes.CDLL(p)
        vp, sz, ci = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int
        L.edison_init.argtypes = [ctypes.POINTER(_Cfg)]
        L.edison_last_error.restype = ctypes.c_char_p
        L.edison_launch_count.restype = ctypes.c_ulonglong
        L.edison_host_alloc.restype = vp
        L.edison_host_alloc.argtypes = [sz]
        L.edison_host_free.argtypes = [vp]
        L.edison_mfcc_q15_batch.argtypes = [vp, sz, vp]
        L.edison_mfcc_q15_clips.argtypes = [vp, sz, sz, sz, vp, vp]
        L.edison_kws_batch.argtypes = [vp, sz, sz, vp, vp, vp]
        L.edison_infer_q7_batch.argtypes = [vp, sz, vp, vp]
        L.edison_mfcc_q15_clips_dev.argtypes = [vp, sz, sz, sz, vp, vp, vp]
        L.edison_infer_q7_batch_dev.argtypes = [vp, sz, vp, vp, vp]
        L.edison_kws_batch_dev.argtypes = [vp, sz, sz, vp, vp, vp, vp]
        L.edison_kws_batch_noalloc_dev.argtypes = [vp, sz, sz, vp, vp, vp]
        L.edison_sync.argtypes = [vp]
        L.edison_kws_counts_dev.argtypes = [vp, sz, sz, vp, vp, vp, vp, vp]
        L.edison_infer_q7_counts_dev.argtypes = [vp, sz, vp, vp, vp, vp]
        L.edison_mfccf_create.argtypes = [ci, ci, ci, ci, ctypes.c_float, ctypes.POINTER(vp)]
        L.edison_mfccf_delete.argtypes = [vp]
        L.edison_mfccf_delete.restype = None
        L.edison_mfccf_n_out.argtypes = [vp]
        L.edison_mfccf_compute_batch.argtypes = [vp, vp, sz, sz, vp]
        L.edison_mfccf_compute_batch_dev.argtypes = [vp, vp, sz, sz, vp, vp]
        L.mfcc_create.argtypes = [ci, ci, ci, ci, ctypes.c_float]
        L.mfcc_create.restype = vp
        L.mfcc_compute.argtypes = [vp, vp, vp]
        L.mfcc_compute.restype = None
        L.mfcc_delete.argtypes = [vp]
        L.mfcc_delete.restype = None
        L.edison_model_load.argtypes = [ctypes.c_char_p, ctypes.POINTER(vp)]
        L.edison_model_free.argtypes = [vp]
        L.edison_model_free.restype = None
        L.edison_model_shape.argtypes = [vp] + [ctypes.POINTER(ci)] * 4
        L.edison_model_infer.argtypes = [vp, vp, sz, vp, vp]
        L.edison_model_set_naive.argtypes = [vp, ci]
        L.edison_infer_f32_batch.argtypes = [vp, sz, vp, vp]
        L.edison_infer_f32_batch_dev.argtypes = [vp, sz, vp, vp, vp]
        for t in ("f32", "f64"):
            getattr(L, "edison_mfcc_%s_batch" % t).argtypes = [vp, sz, vp, ci]
            getattr(L, "edison_mfcc_%s_clips" % t).argtypes = [vp, sz, sz, sz, vp, ci]
            getattr(L, "edison_mfcc_%s_clips_dev" % t).argtypes = [vp, sz, sz, sz, vp, ci, vp]
        L.edison_streams_open.argtypes = [sz, ctypes.POINTER(vp)]
        L.edison_streams_close.argtypes = [vp]
        L.edison_streams_close.restype = None
        L.edison_streams_push_hop.argtypes = [vp, vp, vp, vp, vp]
        L.edison_streams_push_hop_dev.argtypes = [vp, vp, vp, vp, vp, vp]
        L.edison_streams_submit_hop.argtypes = [vp, vp]
        L.edison_streams_collect_hop.argtypes = [vp, vp, vp, vp]
        L.edison_streams_set_incremental.argtypes = [vp, ci]
        L.edison_streams_hops.argtypes = [vp]
        L.edison_streams_hops.restype = ctypes.c_ulonglong
        L.edison_streams_window.argtypes = [vp, sz, vp]
        L.edison_debug_sqrt_q31.argtypes = [vp, sz, vp]
        L.edison_debug_mag_sweep.argtypes = [ctypes.c_uint32, ctypes.c_uint32, vp, vp]
        L.edison_set_softmax_mode.argtypes = [ci]
        L.audioCalcMFCCs.argtypes = [vp, ctypes.POINTER(ctypes.POINTER(ctypes.c_int16))]
        L.aiRunInference.argtypes = [vp, vp]
        L.aiGetKeywordFromIndex.restype = ctypes.c_char_p
        L.aiGetKeywordFromIndex.argtypes = [ctypes.c_uint32]
        L.aiGetKeywordCount.restype = ctypes.c_uint32
        L.edison_net_input_window.restype = ctypes.POINTER(ctypes.c_int8)
        L.aiGetInputShape.argtypes = [ctypes.POINTER(ctypes.c_uint16)] * 2
        L.aiGetInputShape.restype = None
        L.mfccToNetInput.argtypes = [ctypes.POINTER(ctypes.c_int16), ctypes.c_uint16, ctypes.c_uint16, ctypes.c_uint32]
        L.mfccToNetInput.restype = None
        _lib = L
    return _lib


def check(rc, what="edison"):
    if rc != 0:
        raise EdisonError("%s failed (%d): %s" % (what, rc, lib().edison_last_error().decode(errors="replace")))


def init(device=-1, data_dir=None, softmax_mode=SOFTMAX_CMSIS_SIMD, chunk_clips=0):
    """softmax_mode defaults to the firmware build's variant (arm_softmax_q7 SIMD branch), like edison_init(NULL)"""
    cfg = _Cfg(device, data_dir.encode() if data_dir else None, softmax_mode, chunk_clips)
    check(lib().edison_init(ctypes.byref(cfg)), "edison_init")


def shutdown():
    lib().edison_shutdown()


def set_softmax_mode(mode):
    check(lib().edison_set_softmax_mode(mode), "edison_set_softmax_mode")


def launch_count():
    return int(lib().edison_launch_count())


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def mfcc_q15_batch(pcm):
    """pcm: int16 [n_frames, 1024] (or flat) -> int16 [n_frames, 32]   (audioCalcMFCCs per frame)"""
    pcm = np.ascontiguousarray(pcm, dtype=np.int16).reshape(-1)
    if pcm.size % FRAME_LEN:
        raise ValueError("pcm length must be a multiple of 1024")
    n = pcm.size // FRAME_LEN
    out = np.empty((n, N_MEL), dtype=np.int16)
    check(lib().edison_mfcc_q15_batch(_p(pcm), n, _p(out)), "edison_mfcc_q15_batch")
    return out


def mfcc_q15_clips(pcm, n_clips, frames_per_clip, clip_stride, want_mfcc=True, want_net_in=True):
    pcm = np.ascontiguousarray(pcm, dtype=np.int16).reshape(-1)
    need = (n_clips - 1) * clip_stride + frames_per_clip * FRAME_LEN if n_clips else 0
    if pcm.size < need:
        raise ValueError("pcm holds %d samples, need %d" % (pcm.size, need))
    nf = n_clips * frames_per_clip
    mf = np.empty((nf, N_MEL), dtype=np.int16) if want_mfcc else None
    ni = np.empty((nf, N_COEFF), dtype=np.int8) if want_net_in else None
    check(lib().edison_mfcc_q15_clips(_p(pcm), n_clips, frames_per_clip, clip_stride, _p(mf), _p(ni)), "edison_mfcc_q15_clips")
    return mf, ni


def kws_batch(pcm, n_clips, clip_stride=N_FRAMES * FRAME_LEN, want_softmax=True, want_net_in=False):
    """appHifMfccAndInference over n_clips clips -> (logits [n,10], softmax [n,10] | None, net_in [n,403] | None)"""
    pcm = np.ascontiguousarray(pcm, dtype=np.int16).reshape(-1)
    need = (n_clips - 1) * clip_stride + N_FRAMES * FRAME_LEN if n_clips else 0
    if pcm.size < need:
        raise ValueError("pcm holds %d samples, need %d" % (pcm.size, need))
    lg = np.empty((n_clips, N_CLASSES), dtype=np.int8)
    sm = np.empty((n_clips, N_CLASSES), dtype=np.int8) if want_softmax else None
    ni = np.empty((n_clips, NET_IN), dtype=np.int8) if want_net_in else None
    check(lib().edison_kws_batch(_p(pcm), n_clips, clip_stride, _p(lg), _p(sm), _p(ni)), "edison_kws_batch")
    return lg, sm, ni


def infer_q7_batch(net_in, want_softmax=True):
    net_in = np.ascontiguousarray(net_in, dtype=np.int8).reshape(-1, NET_IN)
    n = net_in.shape[0]
    lg = np.empty((n, N_CLASSES), dtype=np.int8)
    sm = np.empty((n, N_CLASSES), dtype=np.int8) if want_softmax else None
    check(lib().edison_infer_q7_batch(_p(net_in), n, _p(lg), _p(sm)), "edison_infer_q7_batch")
    return lg, sm


def infer_f32_batch(net_in):
    """float CNN (Cube.AI build of kws_conv): float32 [n, 403] (MFCC values as floats) -> (logits [n,10], softmax [n,10])"""
    net_in = np.ascontiguousarray(net_in, dtype=np.float32).reshape(-1, NET_IN)
    n = net_in.shape[0]
    lg = np.empty((n, N_CLASSES), dtype=np.float32)
    sm = np.empty((n, N_CLASSES), dtype=np.float32)
    check(lib().edison_infer_f32_batch(_p(net_in), n, _p(lg), _p(sm)), "edison_infer_f32_batch")
    return lg, sm


def mfcc_float_clips(pcm, n_clips, frames_per_clip, clip_stride, mode=MFCC_F32_HOST, dtype=np.float64):
    """float MFCC (mfcc_utils.mfcc / mfcc_mcu of the reference) -> [n_clips*frames_per_clip, 32].
    dtype float64 = the fp64 parity kernel (numpy-equivalent), float32 = the fp32 fast kernel."""
    pcm = np.ascontiguousarray(pcm, dtype=np.int16).reshape(-1)
    need = (n_clips - 1) * clip_stride + frames_per_clip * FRAME_LEN if n_clips and frames_per_clip else 0
    if pcm.size < need:
        raise ValueError("pcm holds %d samples, need %d" % (pcm.size, need))
    dtype = np.dtype(dtype)
    if dtype not in (np.dtype(np.float32), np.dtype(np.float64)):
        raise ValueError("dtype must be float32 or float64")
    out = np.empty((n_clips * frames_per_clip, N_MEL), dtype=dtype)
    fn = lib().edison_mfcc_f64_clips if dtype == np.float64 else lib().edison_mfcc_f32_clips
    check(fn(_p(pcm), n_clips, frames_per_clip, clip_stride, _p(out), mode), "edison_mfcc_float_clips")
    return out


def mfcc_float_batch(pcm, mode=MFCC_F32_HOST, dtype=np.float64):
    """Consecutive 1024-sample frames of pcm; trailing samples that do not fill a frame are ignored, as
    mfcc_utils.mfcc does (frame_count = 1 + (n - 1024) // 1024)."""
    pcm = np.ascontiguousarray(pcm, dtype=np.int16).reshape(-1)
    return mfcc_float_clips(pcm, pcm.size // FRAME_LEN, 1, FRAME_LEN, mode, dtype)


class Model:
    """Any NNoM int8 model of the sequential conv/dense family, from a blob made by tools/convert_nnom_weights.py."""

    def __init__(self, blob_path):
        self._h = ctypes.c_void_p()
        check(lib().edison_model_load(os.fsencode(blob_path), ctypes.byref(self._h)), "edison_model_load")
        d = [ctypes.c_int() for _ in range(4)]
        check(lib().edison_model_shape(self._h, *[ctypes.byref(v) for v in d]), "edison_model_shape")
        self.in_shape = (d[0].value, d[1].value, d[2].value)
        self.in_size = d[0].value * d[1].value * d[2].value
        self.n_out = d[3].value

    def infer(self, x):
        x = np.ascontiguousarray(x, dtype=np.int8).reshape(-1, self.in_size)
        lg = np.empty((x.shape[0], self.n_out), dtype=np.int8)
        sm = np.empty((x.shape[0], self.n_out), dtype=np.int8)
        check(lib().edison_model_infer(self._h, _p(x), x.shape[0], _p(lg), _p(sm)), "edison_model_infer")
        return lg, sm

    def set_naive(self, naive):
        """debug / measurement: True = every node on the one-thread-per-output kernels (no tensor-core nodes)"""
        check(lib().edison_model_set_naive(self._h, 1 if naive else 0), "edison_model_set_naive")

    def close(self):
        if self._h:
            lib().edison_model_free(self._h)
            self._h = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


class MfccC:
    """The firmware's float MFCC in C (firmware/src/audio/mfcc.c): mfcc_create(num, offset, frame_len, dec_bits, preemph).
    compute(pcm, n_frames, hop) -> int8 [n_frames, num - offset]."""

    def __init__(self, num_mfcc=13, offset=1, frame_len=512, dec_bits=8, preemph=0.97):
        self._h = ctypes.c_void_p()
        self.frame_len = frame_len
        check(lib().edison_mfccf_create(num_mfcc, offset, frame_len, dec_bits, preemph, ctypes.byref(self._h)), "edison_mfccf_create")
        self.n_out = int(lib().edison_mfccf_n_out(self._h))

    def compute(self, pcm, n_frames=None, hop=None):
        pcm = np.ascontiguousarray(pcm, dtype=np.int16).reshape(-1)
        hop = self.frame_len if hop is None else hop
        if n_frames is None:
            n_frames = 1 + (pcm.size - self.frame_len) // hop if pcm.size >= self.frame_len else 0
        if n_frames and pcm.size < (n_frames - 1) * hop + self.frame_len:
            raise ValueError("pcm too short for %d frames" % n_frames)
        out = np.empty((n_frames, self.n_out), dtype=np.int8)
        check(lib().edison_mfccf_compute_batch(self._h, _p(pcm), n_frames, hop, _p(out)), "edison_mfccf_compute_batch")
        return out

    def close(self):
        if self._h:
            lib().edison_mfccf_delete(self._h)
            self._h = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


class Streams:
    """n concurrent keyword-spotting streams (appMicMfccInfereContinuous, firmware/src/app.c:288-394).
    push_hop(pcm[n,1024]) -> (softmax int8 [n,10], filt float32 [n,10], detect int32 [n]); detect is the class index when
    the filtered maximum exceeds 0.5, -1 when it does not, -2 while the 31-row window is still filling."""

    def __init__(self, n, incremental=True):
        self.n = int(n)
        self._h = ctypes.c_void_p()
        check(lib().edison_streams_open(self.n, ctypes.byref(self._h)), "edison_streams_open")
        check(lib().edison_streams_set_incremental(self._h, int(incremental)), "edison_streams_set_incremental")

    def close(self):
        if self._h:
            lib().edison_streams_close(self._h)
            self._h = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @property
    def handle(self):
        return self._h

    @property
    def hops(self):
        return int(lib().edison_streams_hops(self._h))

    def push_hop(self, pcm):
        pcm = np.ascontiguousarray(pcm, dtype=np.int16).reshape(-1)
        if pcm.size != self.n * FRAME_LEN:
            raise ValueError("a hop is %d x 1024 samples, got %d" % (self.n, pcm.size))
        sm = np.empty((self.n, N_CLASSES), dtype=np.int8)
        fl = np.empty((self.n, N_CLASSES), dtype=np.float32)
        dt = np.empty(self.n, dtype=np.int32)
        check(lib().edison_streams_push_hop(self._h, _p(pcm), _p(sm), _p(fl), _p(dt)), "edison_streams_push_hop")
        return sm, fl, dt

    def submit_hop(self, pcm):
        """pipelined form: start hop h + 1 while hop h computes (at most two in flight); pcm must stay alive until collected"""
        pcm = np.ascontiguousarray(pcm, dtype=np.int16).reshape(-1)
        if pcm.size != self.n * FRAME_LEN:
            raise ValueError("a hop is %d x 1024 samples, got %d" % (self.n, pcm.size))
        check(lib().edison_streams_submit_hop(self._h, _p(pcm)), "edison_streams_submit_hop")
        return pcm

    def collect_hop(self):
        sm = np.empty((self.n, N_CLASSES), dtype=np.int8)
        fl = np.empty((self.n, N_CLASSES), dtype=np.float32)
        dt = np.empty(self.n, dtype=np.int32)
        check(lib().edison_streams_collect_hop(self._h, _p(sm), _p(fl), _p(dt)), "edison_streams_collect_hop")
        return sm, fl, dt

    def window(self, i):
        w = np.empty(NET_IN, dtype=np.int8)
        check(lib().edison_streams_window(self._h, i, _p(w)), "edison_streams_window")
        return w.reshape(N_FRAMES, N_COEFF)
