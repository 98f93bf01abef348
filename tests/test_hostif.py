"""Wire / ingest formats (edison_b200/hostif.py): the framed UART transfers of
firmware/src/io/hostinterface.c between a host endpoint (what mcu_util.py does) and a device endpoint, NNoM's test_data.bin
container, WAV ingest.  The GPU test serves the two KWS commands from libedison_b200 and checks them against the oracle."""
import os
import socket
import struct
import threading
import wave

import numpy as np
import pytest

import common
from edison_b200 import hostif


def pair():
    a, b = socket.socketpair()
    return a.makefile("rwb", buffering=0), b.makefile("rwb", buffering=0), (a, b)


def test_crc_and_header():
    # hostinterface.c:369-375: 0x1234 + byte sum, modulo 2^16
    assert hostif.crc_sum(b"") == 0x1234
    assert hostif.crc_sum(bytes([255] * 1000)) == (0x1234 + 255000) & 0xFFFF
    x = np.arange(5, dtype=np.int16)
    assert hostif.pack_header(b"<", x, 7) == b"<" + bytes([0x33, 7]) + struct.pack("<I", 5)
    assert [hostif.fmt_byte(t) for t in (np.uint8, np.int8, np.uint16, np.int16, np.uint32, np.int32, np.float32)] == list(range(0x30, 0x37))


@pytest.mark.parametrize("dtype", [np.uint8, np.int8, np.int16, np.uint32, np.float32])
def test_transfer_roundtrip_both_directions(dtype):
    hs, ds, keep = pair()
    host, dev = hostif.HifHost(hs), hostif._Endpoint(ds)
    data = (np.random.default_rng(1).normal(0, 50, 777)).astype(dtype)
    got = {}
    t = threading.Thread(target=lambda: got.update(r=dev._receive(hostif.DELIM_HOST_TO_MCU, dtype)))
    t.start(); host.send_data(data, 0x42); t.join()
    assert got["r"][1] == 0x42 and np.array_equal(got["r"][0], data)
    t = threading.Thread(target=lambda: dev._send(hostif.DELIM_MCU_TO_HOST, data, 9))
    t.start(); back, tag = host.receive_data(); t.join()
    assert tag == 9 and np.array_equal(back, data) and back.dtype == np.dtype(dtype)


def test_wrong_format_and_bad_crc_are_refused():
    hs, ds, keep = pair()
    host, dev = hostif.HifHost(hs), hostif._Endpoint(ds)
    res = {}
    t = threading.Thread(target=lambda: res.update(r=dev._receive(hostif.DELIM_HOST_TO_MCU, np.int16)))
    t.start()
    with pytest.raises(hostif.ProtocolError):
        host.send_data(np.zeros(4, np.int8), 1)            # device wants s16: answers 'F' (hostinterface.c:293-298)
    t.join()
    assert res["r"][0] is None
    # corrupt checksum -> 'C'
    t = threading.Thread(target=lambda: res.update(r=dev._receive(hostif.DELIM_HOST_TO_MCU, np.int8)))
    t.start()
    x = np.arange(10, dtype=np.int8)
    hs.write(hostif.pack_header(b"<", x, 0)); assert hs.read(1) == b"a"
    hs.write(x.tobytes() + struct.pack("<H", (hostif.crc_sum(x.tobytes()) + 1) & 0xFFFF))
    assert hs.read(1) == b"C"
    t.join()
    assert res["r"][0] is None


def test_test_data_bin_container(tmp_path):
    rng = np.random.default_rng(2)
    for n in (5, 128, 300):
        x = rng.integers(-128, 128, (n, 403)).astype(np.int8)
        y = rng.integers(0, 10, n).astype(np.int8)
        p = str(tmp_path / ("t%d.bin" % n))
        hostif.write_test_data_bin(p, x, y)
        assert os.path.getsize(p) == -(-n // 128) * 128 + n * 403      # labels padded per block, data not (nnom_utils.py:95-111)
        x2, y2 = hostif.read_test_data_bin(p, 403)
        assert np.array_equal(x, x2) and np.array_equal(y, y2)


def test_wav_ingest(tmp_path):
    pcm = np.random.default_rng(3).integers(-32768, 32767, 5000).astype(np.int16)
    p = str(tmp_path / "a.wav")
    with wave.open(p, "wb") as w:
        w.setnchannels(1); w.setsampwidth(2); w.setframerate(16000); w.writeframes(pcm.tobytes())
    assert np.array_equal(hostif.read_wav_int16(p), pcm)
    with wave.open(p, "wb") as w:
        w.setnchannels(2); w.setsampwidth(2); w.setframerate(16000); w.writeframes(pcm.tobytes())
    with pytest.raises(ValueError):
        hostif.read_wav_int16(p)


@pytest.mark.gpu
def test_device_serves_kws_commands_from_the_gpu(edison, oracle):
    hs, ds, keep = pair()
    dev = hostif.HifDevice(ds)
    t = threading.Thread(target=dev.serve, daemon=True)
    t.start()
    host = hostif.HifHost(hs)
    pcm = common.synth_frames(31 * 3, seed=12).reshape(3, 31 * 1024)
    want_lg, want_sm, want_ni = oracle.kws(pcm, 3)
    for c in range(3):
        net_in, net_out = host.mfcc_and_infere_on_mcu(pcm[c])            # command 0x4
        assert np.array_equal(net_in.reshape(-1), want_ni[c]) and np.array_equal(net_out, want_sm[c])
        assert np.array_equal(host.infere_on_mcu(want_ni[c]), want_sm[c])  # command 0x3
    assert host.send_command(b"\x00", b"\x00\x10") == hostif.RET_CMD_NOT_FOUND   # microphone commands need the board
    hs.close(); keep[0].close()
    t.join(timeout=5)
