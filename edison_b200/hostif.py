"""Wire and ingest formats of the reference, so its own host tools can drive the GPU library unchanged.

* The UART host interface: firmware/src/io/hostinterface.c.  A transfer is
      '<' (host->MCU) or '>' (MCU->host), format byte '0'..'6' (u8 s8 u16 s16 u32 s32 f32, hostinterface.h:14-20), tag byte,
      u32 little-endian element count; the receiver answers 'a'; then the little-endian payload and a u16 checksum
      (0x1234 + the byte-wise sum, :369-375); the receiver answers '^' (or 'C' on a checksum mismatch, 'F' on a wrong format).
  Commands are single bytes answered with a status byte (0 ok, 1 not found, 2 too few arguments; hifRun :121-176).
  `HifDevice` plays the MCU side on any duplex byte stream (a pty for pyserial, a socket, ...) and serves the KWS commands
  from the GPU: 0x3 aiRunInferenceHif (ai.c:133-183) and 0x4 appHifMfccAndInference (app.c:167-221).
  `HifHost` is the host side as audio/edison/mcu/mcu_util.py implements it (sendCommand :298-323, sendData :234-296,
  receiveData :150-232, waitForMcuReady :325-329), on the same kind of stream.
* WAV files as the reference reads them (16 kHz / 16 bit / mono, audio/edison/mfcc/mfcc.py) and NNoM's `test_data.bin`
  (blocks of 128 labels followed by 128 inputs, the last block's labels zero-padded; nnom_utils.py:67-113).
"""
import struct
import wave

import numpy as np

DELIM_MCU_TO_HOST, DELIM_HOST_TO_MCU = b">", b"<"
DELIM_ACK, DELIM_CRC_FAIL, DELIM_CRC_OK, DELIM_MCU_READY, DELIM_WRONG_DAT_FMT = b"a", b"C", b"^", b"R", b"F"
CRC_SEED = 0x1234
FMT_DTYPES = [np.dtype(t) for t in ("<u1", "<i1", "<u2", "<i2", "<u4", "<i4", "<f4")]     # format byte - 0x30
RET_CMD_OK, RET_CMD_NOT_FOUND, RET_CMD_TOO_FEW_ARGUMENTS = 0, 1, 2
CMD_AI_RUN_INFERENCE, CMD_MFCC_AND_INFERENCE = b"\x03", b"\x04"
TAG_NET_OUT, TAG_NET_INPUT, TAG_NET_OUTPUT = 0x17, 0x20, 0x21


def crc_sum(payload):
    """calcCcrSum, hostinterface.c:369-375"""
    return (CRC_SEED + int(np.frombuffer(payload, dtype=np.uint8).sum(dtype=np.uint64))) & 0xFFFF


def fmt_byte(dtype):
    return 0x30 + FMT_DTYPES.index(np.dtype(dtype).newbyteorder("<"))


def pack_header(direction, data, tag):
    return direction + struct.pack("<BBI", fmt_byte(data.dtype), tag, data.size)


class ProtocolError(RuntimeError):
    pass


class _Endpoint:
    """blocking exact-length reads and writes on a duplex stream exposing read(n)/write(b) (socket.makefile('rwb', 0),
    a pty opened with os.fdopen(fd, 'r+b', 0), serial.Serial, ...)"""

    def __init__(self, stream):
        self.s = stream

    def _read(self, n):
        buf = b""
        while len(buf) < n:
            try:
                chunk = self.s.read(n - len(buf))
            except (ConnectionError, ValueError):     # peer reset / stream closed under us
                chunk = b""
            if not chunk:
                raise EOFError("stream closed")
            buf += chunk
        return buf

    def _write(self, b):
        self.s.write(b)
        if hasattr(self.s, "flush"):
            self.s.flush()

    def _wait_for(self, accepted):
        while True:                                   # waitForByte: skip everything else (hostinterface.c:398-409)
            c = self._read(1)
            if c in accepted:
                return c

    def _send(self, direction, data, tag):
        data = np.ascontiguousarray(data)
        payload = data.astype(data.dtype.newbyteorder("<"), copy=False).tobytes()
        self._write(pack_header(direction, data, tag))
        got = self._wait_for((DELIM_ACK, DELIM_WRONG_DAT_FMT))
        if got != DELIM_ACK:
            raise ProtocolError("receiver rejected the data format")
        self._write(payload + struct.pack("<H", crc_sum(payload)))
        got = self._wait_for((DELIM_CRC_OK, DELIM_CRC_FAIL, DELIM_WRONG_DAT_FMT))
        if got != DELIM_CRC_OK:
            raise ProtocolError("transfer not acknowledged (%r)" % got)

    def _receive(self, direction, want_dtype=None, max_bytes=None):
        self._wait_for((direction,))
        fmt, tag, length = struct.unpack("<BBI", self._read(6))
        if not 0x30 <= fmt <= 0x36:
            raise ProtocolError("invalid format byte 0x%02x" % fmt)
        dt = FMT_DTYPES[fmt - 0x30]
        if want_dtype is not None and dt != np.dtype(want_dtype).newbyteorder("<"):
            self._write(DELIM_WRONG_DAT_FMT)          # hiReceive :293-298
            return None, tag
        self._write(DELIM_ACK)
        nbytes = dt.itemsize * length
        if max_bytes is not None:
            nbytes = min(nbytes, max_bytes)           # hiReceive :305
        payload = self._read(nbytes)
        (crc_in,) = struct.unpack("<H", self._read(2))
        if crc_in != crc_sum(payload):
            self._write(DELIM_CRC_FAIL)
            return None, tag
        self._write(DELIM_CRC_OK)
        return np.frombuffer(payload, dtype=dt).copy(), tag


class HifHost(_Endpoint):
    """the host side of the protocol (what mcu_util.py does over pyserial)"""

    def send_command(self, cmd_byte, args=b""):
        self._write(cmd_byte + args)
        ret = self._wait_for((b"\0", b"\1", b"\2"))
        return ret[0]

    def send_data(self, data, tag):
        self._send(DELIM_HOST_TO_MCU, data, tag)

    def receive_data(self):
        return self._receive(DELIM_MCU_TO_HOST)

    def wait_for_mcu_ready(self):
        self._wait_for((DELIM_MCU_READY,))

    # the two KWS transactions of audio/edison/kws/kws_on_mcu.py
    def infere_on_mcu(self, net_input_q7):
        """kws_on_mcu.py:160-184 (NNoM build): send the int8 net input, get the int8 net output"""
        if self.send_command(CMD_AI_RUN_INFERENCE) != RET_CMD_OK:
            raise ProtocolError("command rejected")
        self.send_data(np.asarray(net_input_q7, dtype=np.int8).reshape(-1), 0)
        out, tag = self.receive_data()
        return out

    def mfcc_and_infere_on_mcu(self, pcm):
        """kws_on_mcu.py:186-216: 31 frames of 1024 int16, one transfer each; returns (net_input [31,13], net_output [10])"""
        pcm = np.asarray(pcm, dtype=np.int16).reshape(31, 1024)
        if self.send_command(CMD_MFCC_AND_INFERENCE) != RET_CMD_OK:
            raise ProtocolError("command rejected")
        for f in range(31):
            self.send_data(pcm[f], f)
            self.wait_for_mcu_ready()
        self.wait_for_mcu_ready()
        net_in, _ = self.receive_data()
        net_out, _ = self.receive_data()
        return net_in.reshape(31, 13), net_out


class HifDevice(_Endpoint):
    """the MCU side, computing on the GPU through the firmware's own entry points of libedison_b200
    (audioCalcMFCCs, mfccToNetInput, aiRunInference); NNoM build (int8 transfers)"""

    def __init__(self, stream):
        super().__init__(stream)
        import ctypes
        from . import capi
        self._ct, self._capi, self._lib = ctypes, capi, capi.lib()
        capi.init()
        self._lib.aiInitialize()

    def _infer(self, net_in):
        ct = self._ct
        x = np.ascontiguousarray(net_in, dtype=np.int8)
        y = np.zeros(10, dtype=np.int8)
        rc = self._lib.aiRunInference(x.ctypes.data_as(ct.c_void_p), y.ctypes.data_as(ct.c_void_p))
        if rc != 0:
            raise ProtocolError("aiRunInference returned %d" % rc)
        return y

    def run_once(self):
        """hifRun (hostinterface.c:121-176) for one command byte.  Returns the command byte served, or None at end of stream."""
        try:
            cmd = self._read(1)
        except EOFError:
            return None
        if cmd == CMD_AI_RUN_INFERENCE:                      # aiRunInferenceHif, ai.c:133-183
            self._write(bytes([RET_CMD_OK]))
            x, _ = self._receive(DELIM_HOST_TO_MCU, np.int8, 403)
            if x is not None:
                if x.size < 403:
                    x = np.concatenate([x, np.zeros(403 - x.size, np.int8)])
                self._send(DELIM_MCU_TO_HOST, self._infer(x[:403]), TAG_NET_OUT)
        elif cmd == CMD_MFCC_AND_INFERENCE:                  # appHifMfccAndInference, app.c:167-221
            self._write(bytes([RET_CMD_OK]))
            ct = self._ct
            x, y = ct.c_uint16(), ct.c_uint16()
            self._lib.aiGetInputShape(ct.byref(x), ct.byref(y))
            for frame in range(y.value):
                pcm, _ = self._receive(DELIM_HOST_TO_MCU, np.int16, 2048)
                buf = np.zeros(1024, dtype=np.int16)
                if pcm is not None:
                    buf[:pcm.size] = pcm[:1024]
                out = ct.POINTER(ct.c_int16)()
                self._lib.audioCalcMFCCs(buf.ctypes.data_as(ct.c_void_p), ct.byref(out))
                self._lib.mfccToNetInput(out, x.value, y.value, frame)
                self._write(DELIM_MCU_READY)
            win = np.ctypeslib.as_array(self._lib.edison_net_input_window(), shape=(403,)).copy()
            net_out = self._infer(win)
            self._write(DELIM_MCU_READY)
            self._send(DELIM_MCU_TO_HOST, win, TAG_NET_INPUT)
            self._send(DELIM_MCU_TO_HOST, net_out, TAG_NET_OUTPUT)
        else:
            self._write(bytes([RET_CMD_NOT_FOUND]))          # everything that needs the board (microphone, LEDs, ...)
        return cmd

    def serve(self):
        try:
            while self.run_once() is not None:
                pass
        except (EOFError, OSError):       # the host hung up in the middle of a transaction
            return


# ---- ingest formats -----------------------------------------------------------------------------------------------------
def read_wav_int16(path, expect_rate=16000):
    """16-bit mono PCM as the reference's tools read it (scipy.io.wavfile in audio/edison/mfcc/mfcc.py:150-160)"""
    with wave.open(path) as w:
        if w.getsampwidth() != 2 or w.getnchannels() != 1:
            raise ValueError("%s: need 16-bit mono PCM" % path)
        if expect_rate and w.getframerate() != expect_rate:
            raise ValueError("%s: sample rate %d, expected %d" % (path, w.getframerate(), expect_rate))
        return np.frombuffer(w.readframes(w.getnframes()), dtype="<i2").astype(np.int16)


def write_test_data_bin(path, x_q7, labels, label_batch=128):
    """generate_test_bin's container (nnom_utils.py:95-113), for inputs that are already q7"""
    x = np.ascontiguousarray(x_q7, dtype=np.int8).reshape(len(labels), -1)
    labels = np.asarray(labels, dtype=np.int8)
    with open(path, "wb") as f:
        for start in range(0, labels.size, label_batch):
            lab = labels[start:start + label_batch]
            f.write(np.pad(lab, (0, label_batch - lab.size)).tobytes())
            f.write(x[start:start + label_batch].tobytes())


def read_test_data_bin(path, input_size, label_batch=128):
    """-> (x int8 [n, input_size], labels int8 [n]); the last block may hold fewer than 128 inputs"""
    raw = np.fromfile(path, dtype=np.int8)
    xs, ls, off = [], [], 0
    while off < raw.size:
        lab = raw[off:off + label_batch]
        off += label_batch
        n = min(label_batch, (raw.size - off) // input_size)
        xs.append(raw[off:off + n * input_size].reshape(n, input_size))
        ls.append(lab[:n])
        off += n * input_size
    return np.concatenate(xs), np.concatenate(ls)
