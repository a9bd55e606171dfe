"""Deterministic, numpy-version-independent synthetic IQ for the golden fixtures: integer-hash noise (splitmix64 of
the sample index, exact uint64 arithmetic) plus complex tones. No RNG distributions are involved."""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix(z):
    with np.errstate(over="ignore"):
        z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
        z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
        z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
        return z ^ (z >> np.uint64(31))


def hash_noise(n, seed):
    """n complex samples, re/im uniform in [-1, 1)."""
    with np.errstate(over="ignore"):
        k = np.arange(n, dtype=np.uint64) * np.uint64(2) + np.uint64(seed) * np.uint64(0x100000001B3)
        a = _splitmix(k)
        b = _splitmix(k + np.uint64(1))
    to_f = lambda u: (u >> np.uint64(11)).astype(np.float64) * (2.0 / 9007199254740992.0) - 1.0
    return to_f(a) + 1j * to_f(b)


def iq(n, seed, fs, tones, noise):
    """noise * hash_noise + sum of amp * exp(j 2 pi f/fs k)"""
    x = hash_noise(n, seed) * noise
    k = np.arange(n, dtype=np.float64)
    for f, amp in tones:
        x = x + amp * np.exp(2j * np.pi * (f / fs) * k)
    return np.ascontiguousarray(x.astype(np.complex128))
