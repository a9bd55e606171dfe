"""GPU test of the one entry point that allocates nothing: pebble_sigstrength_fd_estimate_noalloc_device (include/pebble_cuda.h).

It has no counterpart in the reference (added for callers that cannot call cudaMalloc); its
contract is "the same numbers as pebble_sigstrength_fd_estimate_device", whose own parity against the compiled reference is
tests/test_gpu_prefilter.py::test_fd_estimate_squelch_vs_compiled_reference (1e-9 dB). So the check here is the new entry point
against the existing one through the C ABI, at that same 1e-9 dB, on calls IN A ROW that differ in channel count, band and
spectrum — the scratch array is shared by every call and keeps the last call's contents, and a stale entry would show up as a
wrong channel — plus the size limit. The file sorts last on purpose; it needs neither the
oracle nor any state left by other tests, and passes alone in a fresh process as well as after them."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL_DB = 1e-9


def _spectrum(bins, seed):
    """A dB spectrum like SignalSpectrum's output: noise floor near -100 dB with a few carriers, clipped to [-120, 0]."""
    rng = np.random.default_rng(seed)
    db = -100.0 + 6.0 * rng.standard_normal(bins)
    for c, w, lvl in ((bins // 2 + bins // 8, 9, -30.0), (bins // 2 - bins // 5, 3, -55.0), (bins - 40, 15, -20.0), (17, 5, -70.0)):
        db[c - w:c + w + 1] = np.maximum(db[c - w:c + w + 1], lvl - 0.8 * np.abs(np.arange(-w, w + 1)))
    return np.clip(db, -120.0, 0.0)


def _call(fn, lib, d_spec, bins, fs, lo, hi, mixers, squelch):
    mix = np.ascontiguousarray(mixers, dtype=np.float64)
    out = np.full((len(mix), 5), 12345.0)
    r = fn(d_spec.data_ptr(), bins, int(fs), float(lo), float(hi), mix.ctypes.data, len(mix), float(squelch), out.ctypes.data, None)
    return r, out


def _same(a, b):
    assert a.shape == b.shape
    nan = np.isnan(b)
    assert np.array_equal(np.isnan(a), nan)
    assert np.all(np.abs(a[~nan] - b[~nan]) <= TOL_DB), np.abs(a[~nan] - b[~nan]).max()


def test_fd_estimate_noalloc_matches_allocating_entry_point_on_calls_in_a_row():
    import torch
    from pebblesdr_b200 import lib
    new, old = lib.pebble_sigstrength_fd_estimate_noalloc_device, lib.pebble_sigstrength_fd_estimate_device
    assert lib.pebble_sigstrength_fd_estimate_noalloc_max_channels() == 1024
    fs = 2000000
    cases = [   # (bins, seed, lo, hi, mixers): second call has FEWER channels than the first, third the maximum
        (4096, 1, -5000.0, 5000.0, [2.5e5, -4.0e5, 7.77e5, 0.0, -9.9e5, 9.99e5, 1234.5]),
        (4096, 2, 300.0, 3000.0, [-4.0e5, 9.0e5, 2.5e5]),
        (8192, 3, -100000.0, 100000.0, list(np.linspace(-0.999e6, 0.999e6, 1024))),
        (2048, 4, -5000.0, 5000.0, [7.77e5]),
    ]
    for rep in range(2):                    # the whole sequence twice: the array now holds the 1024-channel call's leftovers
        for bins, seed, lo, hi, mixers in cases:
            d = torch.from_numpy(_spectrum(bins, seed + 10 * rep)).cuda()
            r0, want = _call(old, lib, d, bins, fs, lo, hi, mixers, -60.0)
            r1, got = _call(new, lib, d, bins, fs, lo, hi, mixers, -60.0)
            assert r0 == r1 == len(mixers)
            assert not np.any(got == 12345.0)
            _same(got, want)
            r2, again = _call(new, lib, d, bins, fs, lo, hi, mixers, -60.0)     # immediately again: identical
            assert r2 == len(mixers) and np.array_equal(again, got, equal_nan=True)
    assert want[0, 4] in (0.0, 1.0)


def test_fd_estimate_noalloc_rejects_more_channels_than_the_static_array():
    import torch
    from pebblesdr_b200 import lib
    d = torch.from_numpy(_spectrum(4096, 9)).cuda()
    mixers = list(np.linspace(-9e5, 9e5, 1025))
    r, out = _call(lib.pebble_sigstrength_fd_estimate_noalloc_device, lib, d, 4096, 2000000, -5000.0, 5000.0, mixers, -60.0)
    assert r < 0 and np.all(out == 12345.0)                # refused before touching the output
    msg = lib.pebble_last_error().decode()
    assert "1025" in msg and "1024" in msg and "PEBBLE_FD_NOALLOC_MAX_CH" in msg
    # the allocating entry point takes the same call
    r, out = _call(lib.pebble_sigstrength_fd_estimate_device, lib, d, 4096, 2000000, -5000.0, 5000.0, mixers, -60.0)
    assert r == 1025 and not np.any(out == 12345.0)
    # and the refused call left the no-alloc path usable
    r, out2 = _call(lib.pebble_sigstrength_fd_estimate_noalloc_device, lib, d, 4096, 2000000, -5000.0, 5000.0, mixers[:1024], -60.0)
    assert r == 1024
    _same(out2, out[:1024])
