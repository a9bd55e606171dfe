"""IQ .wav captures in the layout PebbleSDR records and plays (pebblelib/wavfile.cpp:30-160 read, :162-262 write;
plugins/FileSDRDevice/filesdrdevice.cpp:240-256 is the consumer): RIFF/WAVE, `fmt ` (PCM 16-bit or IEEE float, left = I,
right = Q), a `list` chunk of type `info` whose `lofr` / `mode` tags carry the LO frequency and the demodulator mode, `data`.

Harness utility (last row): lets a recorded capture replace the synthetic input of bench.py / the tests. It hands
back the RAW interleaved samples (int16 or float32) — the conversion to CPX is done on the GPU in the mixer's load stage
(pebble_*_set_input_format); the reference scales PCM by 1/32767 (wavfile.cpp:296-297), so `norm_gain(fmt)` gives the
normalise gain that reproduces that scale on top of normalizeIQ's own 1/32768."""
import struct

import numpy as np

PCM_FORMAT, FLOAT_FORMAT = 1, 3


def norm_gain(fmt_tag):
    """Factor for pebble_*_set_input_format's norm_gain so int16 codes are scaled by 1/32767 like WavFile::ReadSamples."""
    return 32768.0 / 32767.0 if fmt_tag == PCM_FORMAT else 1.0


def write_iq_wav(path, raw, sample_rate, lo_freq=0, mode=0):
    """raw: interleaved I,Q int16 (PCM) or float32 (IEEE float). Same chunk order and sizes as WavFile::OpenWrite/Close."""
    raw = np.ascontiguousarray(raw)
    if raw.dtype == np.int16:
        tag, bits = PCM_FORMAT, 16
    elif raw.dtype == np.float32:
        tag, bits = FLOAT_FORMAT, 32
    else:
        raise ValueError("raw must be int16 or float32 interleaved I,Q")
    assert raw.ndim == 1 and raw.size % 2 == 0
    data = raw.tobytes()
    fmt = struct.pack("<HHIIHH", tag, 2, int(sample_rate), int(sample_rate) * 2 * (bits // 8), 2 * (bits // 8), bits)
    # LIST_SUB_CHUNK (wavfile.h:140-146): typeID[4] + text[64], tags as `tag`, uint32 length (incl. NUL), value
    text = b""
    for t, v in ((b"lofr", b"%dl" % int(lo_freq)), (b"mode", b"%d" % int(mode))):      # "%ul" in the reference: digits then a literal 'l'
        v += b"\0"
        text += t + struct.pack("<I", len(v)) + v
    lst = b"info" + text.ljust(64, b"\0")
    body = b"WAVE" + b"fmt " + struct.pack("<I", len(fmt)) + fmt + b"list" + struct.pack("<I", len(lst)) + lst + b"data" + struct.pack("<I", len(data)) + data
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", len(body)) + body)


def read_iq_wav(path):
    """-> dict(raw=interleaved int16 / float32 array (mono files: I only, Q = 0 inserted), sample_rate, lo_freq, mode, fmt_tag,
    channels). Chunk walk as WavFile::OpenRead: ids compared case-insensitively, unknown chunks skipped."""
    with open(path, "rb") as f:
        b = f.read()
    if b[:4] != b"RIFF" or b[8:12] != b"WAVE":
        raise ValueError("not a RIFF/WAVE file")
    pos, info = 12, dict(lo_freq=0, mode=255, sample_rate=0, fmt_tag=0, channels=0, raw=None)
    bits = 16
    while pos + 8 <= len(b):
        cid, size = b[pos:pos + 4].lower(), struct.unpack("<I", b[pos + 4:pos + 8])[0]
        body = b[pos + 8:pos + 8 + size]
        if cid == b"fmt ":
            tag, ch, sr, _, _, bits = struct.unpack("<HHIIHH", body[:16])
            info.update(fmt_tag=tag, channels=ch, sample_rate=sr)
        elif cid == b"list" and body[:4].lower() == b"info":
            i = 4
            while i + 8 <= len(body):
                t, n = body[i:i + 4].lower(), struct.unpack("<I", body[i + 4:i + 8])[0]
                v = body[i + 8:i + 8 + n].split(b"\0")[0]
                i += 8 + n
                if not t.strip(b"\0"):
                    break
                digits = bytes(c for c in v if c in b"0123456789-")
                if t == b"lofr" and digits:
                    info["lo_freq"] = int(digits)
                elif t == b"mode" and digits:
                    info["mode"] = int(digits)
        elif cid == b"data":
            dt = np.int16 if info["fmt_tag"] == PCM_FORMAT else np.float32
            raw = np.frombuffer(body[:len(body) - len(body) % np.dtype(dt).itemsize], dtype=dt)
            if info["channels"] == 1:
                raw = np.stack([raw, np.zeros_like(raw)], axis=1).reshape(-1)
            info["raw"] = np.ascontiguousarray(raw[:raw.size - raw.size % 2])
        pos += 8 + size + (size & 1)
    if info["raw"] is None or not info["sample_rate"]:
        raise ValueError("missing fmt / data chunk")
    if info["fmt_tag"] not in (PCM_FORMAT, FLOAT_FORMAT) or (info["fmt_tag"] == PCM_FORMAT and bits != 16):
        raise ValueError("only 16-bit PCM and IEEE float IQ files are supported (as in the reference)")
    return info
