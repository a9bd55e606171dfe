"""Synthetic RDS for the tests: groups -> 26-bit blocks with checkwords (IEC 62106 annex B: generator polynomial
x^10 + x^8 + x^7 + x^5 + x^4 + x^3 + 1, offset words A/B/C/C'/D) -> differential encoding -> bi-phase impulse pairs ->
the standard's shaping filter H(f) = cos(pi f T / 4) -> DSB on a 57 kHz subcarrier inside an FM stereo multiplex."""
import numpy as np

POLY = 0x5B9
OFFSET = {"A": 0x0FC, "B": 0x198, "C": 0x168, "Cp": 0x350, "D": 0x1B4}
BITRATE = 1187.5
GROUPS = [(0x54A8, 0x0408, 0x2037, 0x4142), (0x54A8, 0x0409, 0x2038, 0x4344), (0x54A8, 0x2410, 0x4845, 0x4C4C),
          (0x54A8, 0x0C0A, 0x54A8, 0x4546)]          # the third is a version-B group (block C carries offset C')


def checkword(msg):
    r = msg << 10
    for b in range(25, 9, -1):
        if (r >> b) & 1:
            r ^= POLY << (b - 10)
    return r & 0x3FF


def block_bits(msg, off):
    w = (msg << 10) | (checkword(msg) ^ OFFSET[off])
    return [(w >> k) & 1 for k in range(25, -1, -1)]


def group_bits(a, b, c, d):
    return block_bits(a, "A") + block_bits(b, "B") + block_bits(c, "Cp" if (b >> 11) & 1 else "C") + block_bits(d, "D")


def stream_bits(groups, nbits):
    bits = []
    while len(bits) < nbits:
        for g in groups:
            bits += group_bits(*g)
    return np.array(bits[:nbits], dtype=np.uint8)


def shaping(fs, span_bits=6):
    T = 1 / BITRATE
    a, B = np.pi * T / 4, 2 / T
    t = np.arange(-int(span_bits * T * fs), int(span_bits * T * fs) + 1) / fs
    h = B * (np.sinc((a + 2 * np.pi * t) * B / np.pi) + np.sinc((a - 2 * np.pi * t) * B / np.pi))
    return h / np.abs(h).max()


def baseband(bits, fs, n):
    from scipy.signal import fftconvolve
    half = fs / (2 * BITRATE)
    assert abs(half - round(half)) < 1e-9, "half a bit must be an integer number of samples"
    half = int(round(half))
    d = np.bitwise_xor.accumulate(np.asarray(bits, dtype=np.int64))          # differential encoding
    imp = np.zeros(n + 4 * half)
    k = np.arange(len(d)) * 2 * half
    ok = k + half < len(imp)
    sym = 2.0 * d - 1.0
    imp[k[ok]] = sym[ok]
    imp[k[ok] + half] = -sym[ok]
    y = fftconvolve(imp, shaping(fs), mode="same")[:n]
    return y / np.abs(y).max()


def fm_broadcast(bits, rate, n, rds_level=0.08, rds_off=0.0, pilot_hz=19000.0, noise=0.001, seed=1):
    """Complex baseband FM signal (75 kHz deviation): L/R tones, pilot, L-R on 38 kHz, RDS on 57 kHz + rds_off."""
    t = np.arange(n) / rate
    left, right = 0.5 * np.sin(2 * np.pi * 1000.0 * t), 0.3 * np.sin(2 * np.pi * 2500.0 * t)
    pil = 2 * np.pi * pilot_hz * t
    m = (0.4 * (left + right) + 0.1 * np.sin(pil) + 0.4 * (left - right) * np.sin(2 * pil) +
         rds_level * baseband(bits, rate, n) * np.cos(2 * np.pi * (57000.0 + rds_off) * t))
    ph = 2 * np.pi * np.cumsum(75000.0 * m) / rate
    rng = np.random.default_rng(seed)
    return np.ascontiguousarray(0.4 * np.exp(1j * ph) + (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * noise)
