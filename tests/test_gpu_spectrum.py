"""GPU parity: FFT operator / SignalSpectrum path vs the reference FFT (FFTW backend) and the C oracle.
Tolerances: forward/inverse rel-RMS <= 1e-10; spectrum <= 1e-6 dB on un-clipped bins, clipped bins
equal; overload flag exact."""
import numpy as np
import pytest

from conftest import rel_rms, iq_noise_tones

pytestmark = pytest.mark.gpu
BH = 12  # WindowFunction::BLACKMANHARRIS


def _spec_input(frames, n, fs, seed=3456):
    # BASELINE config 3: noise 0.01 + 5 tones from -10 to -50 dBFS
    tones = [(-7e6, 10 ** (-10 / 20)), (-1e6, 10 ** (-20 / 20)), (0.3e6, 10 ** (-30 / 20)), (4e6, 10 ** (-40 / 20)),
             (9e6, 10 ** (-50 / 20))]
    return iq_noise_tones(frames * n, seed, fs, tones, noise=0.01)


@pytest.mark.parametrize("size", [2048, 4096, 8192])
def test_forward_inverse_vs_reference(size):
    import pebblesdr_b200 as pb
    from oracle import ref
    g, r = pb.FFT(), ref.FFT()
    g.fftParams(size, 0, 2e6, size, 14)
    r.fftParams(size, 0, 2e6, size, 14)
    x = iq_noise_tones(size, 5, 2e6, [(1e5, 0.5)])
    a, b = np.zeros(size, complex), np.zeros(size, complex)
    g.fftForward(x, a, size)
    r.fftForward(x, b, size)
    assert rel_rms(a, b) <= 1e-10
    a2, b2 = np.zeros(size, complex), np.zeros(size, complex)
    g.fftInverse(b, a2, size)
    r.fftInverse(b, b2, size)
    assert rel_rms(a2, b2) <= 1e-10
    # zero padded short input (fft.cpp:149-153)
    g.fftForward(x[:1000].copy(), a, 1000)
    r.fftForward(x[:1000].copy(), b, 1000)
    assert rel_rms(a, b) <= 1e-10


@pytest.mark.parametrize("size", [2048, 4096, 8192])
def test_spectrum_frames_vs_reference(size):
    import pebblesdr_b200 as pb
    from oracle import ref, port
    frames = 6
    x = _spec_input(frames, size, 20e6)
    x[3 * size + 17] = 0.95 + 0.1j      # one frame with an over-range sample -> overload flag
    g, r, o = pb.FFT(), ref.FFT(), port.FFT()
    for f in (g, r, o):
        f.fftParams(size, 0, 20e6, size, BH)
    for k in range(frames):
        blk = x[k * size:(k + 1) * size].copy()
        dg, dr, do = np.zeros(size), np.zeros(size), np.zeros(size)
        fg, fr, fo = g.fftSpectrum(blk, dg, size), r.fftSpectrum(blk, dr, size), o.fftSpectrum(blk, do, size)
        assert fg == fo == (k == 3)
        assert np.abs(dg - do).max() <= 1e-6          # C oracle (zero-initialised averaging state)
        if k > 0:                                      # the reference's first frame averages with uninitialised memory
            assert fr == fg
            clipped = (dr <= -120.0) | (dr >= 0.0)
            assert np.array_equal(dg[clipped], dr[clipped])
            assert np.abs(dg - dr)[~clipped].max() <= 1e-6
    # -10 dBFS tone reads about -10 dB (worked example at fft.cpp:302-319)
    assert abs(dg.max() - (-10.0)) < 1.0   # scalloping loss depends on the bin offset


def test_spectrum_batch_device_matches_single_calls():
    import torch
    import pebblesdr_b200 as pb
    size, frames = 8192, 40
    x = _spec_input(frames, size, 20e6, seed=9)
    g1, g2 = pb.FFT(), pb.FFT()
    g1.fftParams(size, 0, 20e6, size, BH)
    g2.fftParams(size, 0, 20e6, size, BH)
    single = np.zeros((frames, size))
    for k in range(frames):
        g1.fftSpectrum(x[k * size:(k + 1) * size].copy(), single[k], size)
    xd = torch.from_numpy(x).cuda()
    out = torch.zeros(frames, size, dtype=torch.float64, device="cuda")
    flags = torch.zeros(frames, dtype=torch.int32, device="cuda")
    # two batches: the carried averaging state must cross the call boundary
    assert g2.fftSpectrumBatch(25, size, xd[:25 * size], out[:25], flags[:25]) == 25
    assert g2.fftSpectrumBatch(15, size, xd[25 * size:], out[25:], flags[25:]) == 15
    assert np.abs(out.cpu().numpy() - single).max() <= 1e-9
    assert int(flags.sum().item()) == 0


def test_window_tables_bit_exact():
    import pebblesdr_b200 as pb
    from oracle import ref
    for t in range(1, 13):
        for n in (2048, 8192, 1000):
            w = pb.WindowFunction(t, n)
            rw, rcg = ref.window(t, n)
            assert np.array_equal(w.window, rw) and w.coherentGain == rcg


def test_spectrum_db_edges_zero_tiny_and_full_scale_frames():
    """The table-reduced dB conversion (spec_db) against the oracle on the edges of its range: an all-zero frame (bin == 0
    -> -120 exactly, db.h:44-48), amplitudes whose spectrum clips at -120, a wide dynamic range (1e-7 .. 0.9 across the
    band) and a full-scale tone that sets the overload flag. Clipped bins equal, others within 1e-6 dB, flags exact."""
    import pebblesdr_b200 as pb
    from oracle import port
    size = 8192
    g, o = pb.FFT(), port.FFT()
    g.fftParams(size, 0, 20e6, size, 12)
    o.fftParams(size, 0, 20e6, size, 12)
    t = np.arange(size)
    rng = np.random.default_rng(99)
    frames = [
        np.zeros(size, complex),
        1e-9 * np.exp(2j * np.pi * 0.11 * t),
        sum(10.0 ** (-k) * np.exp(2j * np.pi * (0.03 + 0.06 * k) * t) for k in range(1, 8)) + 1e-7 * rng.standard_normal(size),
        0.95 * np.exp(2j * np.pi * (1024 / size) * t),
        np.zeros(size, complex),
    ]
    for f, x in enumerate(frames):
        x = np.ascontiguousarray(x.astype(np.complex128))
        yg, yo = np.zeros(size), np.zeros(size)
        fg, fo = g.fftSpectrum(x, yg, size), o.fftSpectrum(x, yo, size)
        assert bool(fg) == bool(fo), f
        clipped = (yo <= -120.0) | (yo >= 0.0)
        assert np.array_equal(yg[clipped], yo[clipped]), f
        assert np.max(np.abs(yg[~clipped] - yo[~clipped]), initial=0.0) <= 1e-6, f
    # last frame: zeros after a loud frame -> (0 + previous UNAVERAGED amplitude) / 2 (F14), not all -120
    assert yg.max() > -120.0 and np.count_nonzero(yg == -120.0) > size // 2


def test_ooura_backend_contract_vs_compiled_cfft():
    """class CFft (pebblelib/fftcute.cpp, the -DUSE_FFTCUTE build of oracle/build_ref.sh step 7). fftForward swaps I
    and Q before Ooura's transform (:194-201), fftInverse is Ooura's e^{-j} 'backward' transform, fftSpectrum equals FFTW's."""
    import pebblesdr_b200 as pb
    from oracle import ref_app
    if not ref_app.cute_available():
        pytest.skip("oracle/_ref/libpebbleref_cute.so not built")
    for N, wtype in ((2048, 12), (4096, 12), (8192, 6)):
        x = iq_noise_tones(N, 70 + N, 2e6, [(1.23e5, 0.4), (-3e5, 0.2)], noise=0.05)
        c = ref_app.CFft(); c.fftParams(N, 0, 2e6, N, wtype)
        g = pb.FFT(); g.fftParams(N, 0, 2e6, N, wtype); g.setBackend(1)
        a, b = np.zeros(N, dtype=np.complex128), np.zeros(N, dtype=np.complex128)
        c.fftForward(x, a, N); g.fftForward(x, b, N)
        assert rel_rms(b, a) <= 1e-12
        xi = iq_noise_tones(N, 80 + N, 2e6, [(5e4, 0.3)], noise=0.1)
        c.fftInverse(xi, a, N); g.fftInverse(xi, b, N)
        assert rel_rms(b, a) <= 1e-12
        # and the FFTW contract differs from it exactly by the I/Q exchange
        g2 = pb.FFT(); g2.fftParams(N, 0, 2e6, N, wtype)
        b2 = np.zeros(N, dtype=np.complex128)
        g2.fftForward(x, b2, N); g.fftForward(x, b, N)
        assert np.array_equal(b, b2.imag + 1j * b2.real)
        da, db = np.zeros(N), np.zeros(N)
        for _ in range(3):
            fa = c.fftSpectrum(x, da, N); fb = g.fftSpectrum(x, db, N)
        un = (da > -120) & (da < 0)
        assert un.any() and fa == fb and np.max(np.abs(da[un] - db[un])) <= 1e-6 and np.array_equal(da[~un], db[~un])


def test_signal_spectrum_operator_gate_and_two_ffts_vs_compiled_reference():
    """SignalSpectrum::unprocessed / zoomed / setUpdatesPerSec / setSampleRate (signalspectrum.cpp:54-133) against
    the class itself compiled from the reference with a settable clock in place of QElapsedTimer (oracle/qt_shim_app)."""
    import pebblesdr_b200 as pb
    from oracle import ref_app
    fs, hfs, n = 2000000, 62500, 2048
    ref_app.set_clock_ms(0)
    r = ref_app.SignalSpectrum(fs, hfs, n, bins=4096, hires_bins=2048, updates_per_sec=10)
    g = pb.SignalSpectrum(fs, hfs, n, 4096, 2048, 10)
    assert g.binCount() == 4096
    r.setHiRes(True); g.setHiRes(True)
    made = 0
    # 1 ms steps would take long: walk the clock in uneven steps around the 100 ms gate, change the rate mid-way
    clock = [0, 5, 99, 100, 150, 199, 200, 201, 330, 331, 400, 433, 434, 500, 533, 534, 535, 600, 635, 1000, 1001, 1100]
    for i, t in enumerate(clock):
        if i == 9:
            r.setUpdatesPerSec(30); g.setUpdatesPerSec(30)          # 1000 / 30 = 33 ms (integer division, :129-130)
        if i == 17:
            r.setUpdatesPerSec(0); g.setUpdatesPerSec(0)            # 0: no more updates, interval kept (:126-131)
        if i == 19:
            r.setUpdatesPerSec(10); g.setUpdatesPerSec(10)
        x = iq_noise_tones(n, 500 + i, fs, [(2e5 + 1e3 * i, 0.3), (-6e5, 0.05)], noise=0.01)
        if i == 12:
            x[7] = 0.95                                             # trips the overload flag (fft.cpp:138)
        z = iq_noise_tones(n, 600 + i, hfs, [(1e3, 0.3)], noise=0.01)
        ref_app.set_clock_ms(t)
        before, ov_before = r.updatePending(), r.overrun()
        r.unprocessed(x, n); r.zoomed(z, n)
        new_ref = (r.updatePending() and not before) or r.overrun() > ov_before
        new_g = g.unprocessed(x, n, t)
        g.zoomed(z, n, t)
        assert bool(new_g) == bool(new_ref), (i, t)
        if new_g:
            made += 1
            a, b = r.getUnprocessed(), g.getUnprocessed()
            if made > 1:      # the first spectrum averages with the reference's uninitialised buffer (fft.cpp:348-397)
                un = (a > -120) & (a < 0)
                assert np.max(np.abs(a[un] - b[un])) <= 1e-6 and np.array_equal(a[~un], b[~un])
            assert r.getOverload() == g.getOverload()
            if made % 3:
                r.ackUpdate(); g.ackUpdate()                        # the widget marks the update consumed (else: an overrun next time)
        assert r.overrun() == g.overrun()
    assert made >= 8 and g.overrun() >= 2
