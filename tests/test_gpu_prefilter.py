"""GPU parity, SURVEY 8f-3: CIir biquads / DCRemoval, IQBalance and the squelch estimate SignalStrength::fdEstimate against the
COMPILED reference (application/{dcremoval,iqbalance,signalstrength}.cpp + pebblelib/iir.cpp, oracle/build_ref.sh step 5).

Tolerances. IQBalance: bit-identical (same operations in the same order). fdEstimate: 1e-9 dB (tree-ordered sums). CIir:
1e-10 relative RMS for ordinary filters; for the DC blocker at MHz rates the reference's own direct-form-2 arithmetic keeps
only ~8 digits (its state w is ~1e8 x the signal and y is w's second difference), which test_dc_removal... demonstrates against
an 80-bit evaluation before comparing at the precision the reference actually has."""
import numpy as np
import pytest

from conftest import rel_rms, iq_noise_tones

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind,f0,q,fs", [(0, 75000.0, 1.0, 250000.0), (2, 19000.0, 500.0, 250000.0), (3, 19000.0, 5.0, 250000.0),
                                          (1, 300.0, 0.7071, 48000.0)])
def test_ciir_vs_compiled_reference_ragged_calls(kind, f0, q, fs):
    import pebblesdr_b200 as pb
    from oracle import ref_app
    n = 4096 * 9 + 777
    x = iq_noise_tones(n, 5 + kind, fs, [(f0 * 0.9, 0.4), (-f0 * 1.7 % (fs / 2), 0.3)], noise=0.1)
    r = ref_app.CIir(kind, f0, q, fs)
    want = r.ProcessFilter(x)
    g = pb.CIir(kind, f0, q, fs)
    outs, pos = [], 0
    for m in (1, 15, 4096, 4097, 8192 * 3, n - 1 - 15 - 4096 - 4097 - 8192 * 3):
        o = np.zeros(m, dtype=np.complex128)
        assert g.ProcessFilter(m, np.ascontiguousarray(x[pos:pos + m]), o) == m
        outs.append(o); pos += m
    assert pos == n
    assert rel_rms(np.concatenate(outs), want) <= 1e-10


def test_dc_removal_device_rate_and_the_reference_own_precision():
    """DCRemoval (10 Hz high-pass at 10 MSPS) over 2^22 samples with a DC offset."""
    import torch
    import pebblesdr_b200 as pb
    from oracle import ref_app
    fs, n = 10000000, 1 << 22
    x = iq_noise_tones(n, 77, fs, [(1.2e6, 0.3)], noise=0.05) + (0.2 - 0.1j)
    r = ref_app.DCRemoval(fs, 65536)        # the step owns an `out` of bufferSize samples
    r.enableStep(True)
    want = np.concatenate([r.process(x[i:i + 65536]) for i in range(0, n, 65536)])
    g = pb.DCRemoval(fs)
    g.enableStep(True)
    out = torch.zeros(n, dtype=torch.complex128, device="cuda")
    half = n // 2 + 4096 * 3 + 5
    xd = torch.from_numpy(x).cuda()
    g.ProcessFilter(half, xd[:half], out[:half])
    g.ProcessFilter(n - half, xd[half:], out[half:])
    got = out.cpu().numpy()
    # what the reference's arithmetic is worth: the same recurrence in 80-bit on a stretch of the stream
    c = g.coefficients()
    m = 1 << 16
    a1, a2, b0, b1, b2 = [np.longdouble(v) for v in c]
    w1 = w2 = np.longdouble(0)
    exact = np.zeros(m)
    xr = x.real[:m].astype(np.longdouble)
    for i in range(m):
        w0 = xr[i] - a1 * w1 - a2 * w2
        exact[i] = float(b0 * w0 + b1 * w1 + b2 * w2)
        w2, w1 = w1, w0
    ref_err = rel_rms(want.real[:m], exact)
    got_err = rel_rms(got.real[:m], exact)
    assert got_err <= max(10 * ref_err, 1e-12), (got_err, ref_err)
    # and the two double evaluations agree to the precision either of them has
    assert rel_rms(got, want) <= max(100 * ref_err, 1e-9), (rel_rms(got, want), ref_err)
    # the DC offset is gone, the tone is not
    assert abs(got[n // 2:].mean()) < 1e-3 and np.abs(got[n // 2:]).mean() > 0.2
    # a disabled step copies (dcremoval.cpp:11-14)
    g2 = pb.DCRemoval(fs)
    o = np.zeros(64, dtype=np.complex128)
    g2.process(x[:64], 64, o)
    assert np.array_equal(o, x[:64])


def test_iqbalance_bit_exact():
    import torch
    import pebblesdr_b200 as pb
    from oracle import ref_app
    n, blk = 2048 * 64, 2048
    x = iq_noise_tones(n, 9, 2e6, [(2e5, 0.5), (-2e5, 0.05)], noise=0.02)
    r = ref_app.IQBalance(2000000, blk)
    for gain, phase in ((1.0, 0.0), (1.07, -0.03)):
        r.config(True, gain, phase)
        want = np.concatenate([r.ProcessBlock(x[i:i + blk])[0] for i in range(0, n, blk)])
        got = pb.iqbalance(torch.from_numpy(x).cuda(), blk, gain, phase).cpu().numpy()
        assert np.array_equal(got, want), rel_rms(got, want)


def test_fd_estimate_squelch_vs_compiled_reference():
    """The spectrum is the GPU's own SignalSpectrum output (4096 bins, 2 MSPS); fdEstimate per channel vs the reference class."""
    import torch
    import pebblesdr_b200 as pb
    from oracle import ref_app
    fs, bins, n = 2000000, 4096, 2048
    x = iq_noise_tones(n, 21, fs, [(2.5e5, 0.3), (-4.0e5, 0.01), (7.77e5, 0.1)], noise=0.002)
    f = pb.FFT(); f.fftParams(bins, 0, fs, n, 12)
    db = np.zeros(bins)
    for _ in range(2):
        f.fftSpectrum(x, db, n)
    mixers = [2.5e5, -4.0e5, 7.77e5, 0.0, -9.9e5, 9.99e5, 1234.5]
    for lo, hi in ((-5000.0, 5000.0), (300.0, 3000.0), (-100000.0, 100000.0)):
        got = pb.fdEstimate(torch.from_numpy(db).cuda(), fs, lo, hi, mixers, squelchDb=-60.0)
        for k, mf in enumerate(mixers):
            ref_app.set_clock_ms(0)
            s = ref_app.SignalStrength(fs, n)
            s.fdEstimate(db, bins, fs, lo, hi, mf)      # the first call only starts its 100 ms display timer (:292-300)
            ref_app.set_clock_ms(1000)
            avg = s.fdEstimate(db, bins, fs, lo, hi, mf)
            v = s.values()                              # peakDb, avgDb, snrDb, floorDb, extValue
            assert abs(avg - v[1]) == 0
            for j in range(4):
                if np.isnan(v[j]):
                    assert np.isnan(got[k, j])
                else:
                    assert abs(got[k, j] - v[j]) <= 1e-9, (lo, hi, mf, j, got[k], v)
            assert got[k, 4] == (0.0 if v[1] < -60.0 else 1.0)
    assert got[:, 4].min() == 0.0 and pb.fdEstimate(torch.from_numpy(db).cuda(), fs, -5000.0, 5000.0, mixers[:3], -60.0)[0, 4] == 1.0


def _impulsive(n, seed):
    """Noise with lightning-like spikes (what NB1 blanks) and bursts (what NB2 replaces), plus a carrier."""
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * 0.05 + 0.1 * np.exp(2j * np.pi * 0.0123 * np.arange(n))
    for p in rng.integers(100, n - 100, size=n // 700):
        x[p:p + int(rng.integers(1, 4))] *= 12.0
    return np.ascontiguousarray(x)


def test_noise_blankers_bit_identical_to_compiled_reference():
    """NoiseBlanker::ProcessBlock / ProcessBlock2 (noiseblanker.cpp:47-98) against the class compiled from the reference,
    block after block (the average, the spike run and the two-sample delay line cross block boundaries). Outputs are copies of
    inputs, zeros or the running average, and every decision is taken on the reference's own float arithmetic: np.array_equal."""
    import pebblesdr_b200 as pb
    from oracle import ref_app
    fs, frame, frames = 2000000, 2048, 40
    x = _impulsive(frame * frames, 3)
    r, g = ref_app.NoiseBlanker(fs, frame), pb.NoiseBlanker(fs, frame)
    # disabled (the default): both return the input
    b = x[:frame]
    assert g.ProcessBlock(b) is b and g.ProcessBlock2(b) is b and np.array_equal(r.ProcessBlock(b)[0], b)
    r.enable(True, True); g.setNbEnabled(True); g.setNb2Enabled(True)
    blanked = replaced = 0
    for i in range(0, len(x), frame):
        b = np.ascontiguousarray(x[i:i + frame])
        w1, g1 = r.ProcessBlock(b)[0], g.ProcessBlock(b)
        assert np.array_equal(g1, w1), (i, int(np.sum(g1 != w1)))
        w2, g2 = r.ProcessBlock2(w1)[0], g.ProcessBlock2(g1)         # chained as receiver.cpp:821-822 does
        assert np.array_equal(g2, w2), (i, int(np.sum(g2 != w2)))
        blanked += int(np.sum(w1 == 0)); replaced += int(np.sum(w2 != w1))
    assert blanked > 100 and replaced > 100                        # both blankers actually fired
    # re-enabling restarts the averages (setNbEnabled / setNb2Enabled), ragged block sizes
    r.enable(True, True); g.setNbEnabled(True); g.setNb2Enabled(True)
    r2 = ref_app.NoiseBlanker(fs, 777); r2.enable(True, True)
    g3 = pb.NoiseBlanker(fs, 777); g3.setNbEnabled(True); g3.setNb2Enabled(True)
    for i in range(0, 777 * 5, 777):
        b = np.ascontiguousarray(x[i:i + 777])
        assert np.array_equal(g3.ProcessBlock(b), r2.ProcessBlock(b)[0])
        assert np.array_equal(g3.ProcessBlock2(b), r2.ProcessBlock2(b)[0])
