"""CPU test of the IQ .wav harness reader / writer (pebblesdr_b200/wavfile.py; layout of pebblelib/wavfile.cpp:30-262)."""
import struct

import numpy as np

from pebblesdr_b200 import wavfile


def test_iq_wav_round_trip_pcm_and_float(tmp_path):
    rng = np.random.default_rng(1)
    raw16 = rng.integers(-32768, 32767, size=2 * 5000, dtype=np.int16)
    p = tmp_path / "cap.wav"
    wavfile.write_iq_wav(str(p), raw16, 2000000, lo_freq=14070000, mode=7)
    w = wavfile.read_iq_wav(str(p))
    assert np.array_equal(w["raw"], raw16) and w["sample_rate"] == 2000000 and w["lo_freq"] == 14070000 and w["mode"] == 7
    assert w["fmt_tag"] == wavfile.PCM_FORMAT and w["channels"] == 2 and wavfile.norm_gain(w["fmt_tag"]) == 32768.0 / 32767.0
    b = p.read_bytes()
    # chunk layout the reference writes: RIFF size, fmt (16 bytes), `list` of sizeof(LIST_SUB_CHUNK) = 68 bytes, data
    assert b[:4] == b"RIFF" and struct.unpack("<I", b[4:8])[0] == len(b) - 8 and b[12:16] == b"fmt " and b[36:40] == b"list"
    assert struct.unpack("<I", b[40:44])[0] == 68 and b[44:48] == b"info" and b[48:52] == b"lofr" and b[112:116] == b"data"
    rawf = rng.standard_normal(2 * 777).astype(np.float32)
    wavfile.write_iq_wav(str(p), rawf, 48000)
    w = wavfile.read_iq_wav(str(p))
    assert np.array_equal(w["raw"], rawf) and w["fmt_tag"] == wavfile.FLOAT_FORMAT and w["lo_freq"] == 0 and w["mode"] == 0


def test_iq_wav_skips_unknown_chunks_and_reads_mono(tmp_path):
    mono = np.arange(-50, 50, dtype=np.int16)
    fmt = struct.pack("<HHIIHH", 1, 1, 8000, 16000, 2, 16)
    body = (b"WAVE" + b"JUNK" + struct.pack("<I", 3) + b"abc\0" + b"fmt " + struct.pack("<I", 16) + fmt +
            b"data" + struct.pack("<I", mono.nbytes) + mono.tobytes())
    p = tmp_path / "mono.wav"
    p.write_bytes(b"RIFF" + struct.pack("<I", len(body)) + body)
    w = wavfile.read_iq_wav(str(p))
    assert w["channels"] == 1 and np.array_equal(w["raw"][0::2], mono) and not w["raw"][1::2].any() and w["mode"] == 255
