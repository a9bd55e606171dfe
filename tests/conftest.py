import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def rel_rms(a, b):
    """relative RMS error of a against reference b"""
    a = np.asarray(a)
    b = np.asarray(b)
    den = np.sqrt(np.mean(np.abs(b) ** 2))
    return float(np.sqrt(np.mean(np.abs(a - b) ** 2)) / (den if den > 0 else 1.0))


def iq_noise_tones(n, seed, fs, tones, noise=0.1):
    """Seeded synthetic IQ: complex Gaussian noise + complex tones [(freq_hz, amplitude), ...]."""
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * noise
    t = np.arange(n)
    for f, a in tones:
        x += a * np.exp(2j * np.pi * (f / fs) * t)
    return np.ascontiguousarray(x.astype(np.complex128))
