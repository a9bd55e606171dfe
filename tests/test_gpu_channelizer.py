"""GPU parity: the batched channelizer (BASELINE config 4 shape: 61.44 MSPS capture, 10-stage chain to 60 kHz,
CFastFIR +-7.5 kHz, even channels NBFM PLL, odd channels AM) against the C oracle's Receiver run once per channel.
Tolerances: AM / pass-through channels rel-RMS <= 1e-10, NBFM PLL (float state) <= 1e-4; sample counts exact.
The reference itself cannot run this chain (MAX_DECSTAGES overflow, finding F8); the oracle lifts the limit."""
import numpy as np
import pytest

from conftest import rel_rms

pytestmark = pytest.mark.gpu
FS, BW = 61.44e6, 15000.0


def _capture(n, freqs, seed=4567):
    rng = np.random.default_rng(seed)
    t = np.arange(n, dtype=np.float64)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * 0.002
    for k, f in enumerate(freqs):
        if k % 2 == 0:   # NBFM: 5 kHz deviation, 1 kHz tone
            ph = 2 * np.pi * (f / FS) * t + (5000.0 / 1000.0) * np.sin(2 * np.pi * (1000.0 / FS) * t)
            x += (1.0 / 16) * np.exp(1j * ph)
        else:            # AM: 1 kHz tone, m = 0.5
            x += (1.0 / 16) * (1 + 0.5 * np.cos(2 * np.pi * (1000.0 / FS) * t)) * np.exp(2j * np.pi * (f / FS) * t)
    return np.ascontiguousarray(x.astype(np.complex128))


def _oracle_channel(x, f, mode):
    from oracle import port
    o = port.Receiver(FS, 2048, BW)
    assert o.demod_rate == 60000.0
    o.config(f, -7500.0, 7500.0, mode, 0, 20)
    return np.concatenate([o.processIQData(x[i:i + 65536], 65536) for i in range(0, len(x), 65536)])


def test_channelizer_vs_oracle_per_channel():
    import torch
    import pebblesdr_b200 as pb
    nch = 6
    n = 1024 * 2048 * 2
    freqs = [-30.69e6 + 60e3 * (k * 170 + 3) for k in range(nch)]
    modes = [2, 0, 2, 0, 7, 0]        # dmFMN, dmAM, ..., one dmUSB pass-through channel
    x = _capture(n, freqs)
    cz = pb.Channelizer(nch, FS, BW)
    assert cz.outRate == 60000.0 and cz.numStages == 10
    cz.setFilter(-7500.0, 7500.0)
    for k in range(nch):
        cz.setFrequency(k, freqs[k])
        cz.setDemodMode(k, modes[k])
    xd = torch.from_numpy(x).cuda()
    nd = n >> 10
    out = torch.zeros(nch, nd, dtype=torch.complex128, device="cuda")
    # two calls: carried state (delay lines, NCO phase, overlap, PLL, DC block, CFir history) must cross the boundary
    half = n // 2
    assert cz.process(xd[:half], half, out[:, :nd // 2]) == nd // 2
    o2 = torch.zeros(nch, nd // 2, dtype=torch.complex128, device="cuda")
    assert cz.process(xd[half:], half, o2) == nd // 2
    got = np.concatenate([out[:, :nd // 2].cpu().numpy(), o2.cpu().numpy()], axis=1)
    for k in range(nch):
        want = _oracle_channel(x, freqs[k], modes[k])
        assert len(want) == nd
        if modes[k] == 2:
            # the PLL (float state) is chaotic until it locks: 1e-14 input differences decide the acquisition path,
            # so the NBFM channels are compared after lock (second half of the stream)
            assert rel_rms(got[k][nd // 2:], want[nd // 2:]) <= 1e-4, (k, rel_rms(got[k][nd // 2:], want[nd // 2:]))
        else:
            assert rel_rms(got[k], want) <= 1e-10, (k, modes[k], rel_rms(got[k], want))


def test_batched_downconvert_equals_single_channel_objects():
    import torch
    import pebblesdr_b200 as pb
    nch, n = 5, 1024 * 512
    freqs = [1.2e6, -7.5e6, 20.001e6, 0.0, -30.0e6]
    x = _capture(n, freqs[:2], seed=99)
    xd = torch.from_numpy(x).cuda()
    cz = pb.Channelizer(nch, FS, BW)
    for k, f in enumerate(freqs):
        cz.setFrequency(k, f)
    out = torch.zeros(nch, n >> 10, dtype=torch.complex128, device="cuda")
    assert cz.downConvert(xd, n, out) == n >> 10
    for k, f in enumerate(freqs):
        dc = pb.CDownConvert()
        assert dc.SetDataRate(FS, BW) == 60000.0
        dc.SetFrequency(f)
        one = torch.zeros(n >> 10, dtype=torch.complex128, device="cuda")
        assert dc.ProcessData(n, xd, one) == n >> 10
        assert rel_rms(out[k].cpu().numpy(), one.cpu().numpy()) <= 1e-13


# ---- polyphase-DFT front end (channels on the grid fs / 2^S) ----------------------------------------------------------------
GRID0, STEP = -30.69e6, 60e3       # BASELINE config 4: centre k = -30.69 MHz + k * 60 kHz, and 61.44 MHz / 1024 = 60 kHz


def _ref12_channel(x, f):
    """The COMPILED reference CDownConvert (MAX_DECSTAGES 12, oracle/build_ref.sh step 6) on one channel. 65536-sample calls:
    the 35-tap last stage must receive >= 2 * 34 samples per call or the reference corrupts its own carried history
    (downconvert.cpp:387-390, tests/test_gpu_downconvert.py::test_short_call_rule...)."""
    from oracle import ref_app
    r = ref_app.CDownConvert12()
    assert r.SetDataRate(FS, BW) == 60000.0
    r.SetFrequency(f)
    out = []
    for i in range(0, len(x), 65536):
        o = np.zeros(64, dtype=np.complex128)
        assert r.ProcessData(65536, x[i:i + 65536], o) == 64
        out.append(o)
    return np.concatenate(out)


def test_polyphase_front_end_vs_compiled_reference_per_channel():
    """Grid channels: ONE mix + 1024-branch polyphase FIR + 1024-point DFT (chz_poly.cu) against the reference's own
    CDownConvert run once per channel. 2^22 capture samples in ragged calls; every channel <= 1e-10 relative RMS."""
    import torch
    import pebblesdr_b200 as pb
    from oracle import ref_app
    if not ref_app.dec12_available():
        pytest.skip("oracle/_ref/libpebbleref_dec12.so not built")
    ks = [0, 1, 2, 255, 511, 512, 513, 777, 1022, 1023]
    freqs = [GRID0 + STEP * k for k in ks]
    n = 1 << 22
    x = _capture(n, freqs, seed=123)
    xd = torch.from_numpy(x).cuda()
    cz = pb.Channelizer(len(ks), FS, BW)
    for i, f in enumerate(freqs):
        cz.setFrequency(i, f)
    nd = n >> 10
    out = torch.zeros(len(ks), nd, dtype=torch.complex128, device="cuda")
    pos = 0
    for m in [1024 * 7, 1024 * 1, 1024 * 300, 1024 * 24, 1024 * 1000, n - 1024 * 1332]:   # shorter and longer than the filter memory
        assert cz.downConvert(xd[pos:pos + m], m, out[:, pos >> 10:]) == m >> 10
        pos += m
    assert cz.frontPath() == 1
    got = out.cpu().numpy()
    for i, f in enumerate(freqs):
        want = _ref12_channel(x, f)
        assert rel_rms(got[i], want) <= 1e-10, (ks[i], rel_rms(got[i], want))


def test_polyphase_equals_per_channel_kernel_and_raw_formats(monkeypatch):
    """Same capture through the polyphase path and (PEBBLE_CHZ_DIRECT=1) the per-channel kernel; then the capture as
    qint16 / quint8 raw samples through the polyphase path vs the converted capture."""
    import torch
    import pebblesdr_b200 as pb
    ks = [3, 4, 100, 640, 1000]
    freqs = [GRID0 + STEP * k for k in ks]
    n = 1 << 21
    x = _capture(n, freqs, seed=321)
    xd = torch.from_numpy(x).cuda()

    def run(fmt=None, src=None):
        cz = pb.Channelizer(len(ks), FS, BW)
        for i, f in enumerate(freqs):
            cz.setFrequency(i, f)
        if fmt is not None:
            cz.setInputFormat(fmt)
        o = torch.zeros(len(ks), n >> 10, dtype=torch.complex128, device="cuda")
        half = n // 2
        s = xd if src is None else src
        per = 1 if fmt in (None, 0) else 2
        assert cz.downConvert(s[:per * half], half, o[:, :half >> 10]) == half >> 10
        assert cz.downConvert(s[per * half:], half, o[:, half >> 10:]) == half >> 10
        return cz.frontPath(), o.cpu().numpy()

    path, poly = run()
    assert path == 1
    monkeypatch.setenv("PEBBLE_CHZ_DIRECT", "1")
    path, direct = run()
    assert path == 0
    monkeypatch.delenv("PEBBLE_CHZ_DIRECT")
    for i in range(len(ks)):
        assert rel_rms(poly[i], direct[i]) <= 1e-10, (ks[i], rel_rms(poly[i], direct[i]))
    f2 = np.stack([x.real, x.imag], axis=1).reshape(-1)
    for fmt, raw, scale in ((pb.IQ_CS16, np.round(f2 * 30000).astype(np.int16), 1 / 32768.0),
                            (pb.IQ_CU8, (np.round(f2 * 120) + 128).astype(np.uint8), 1 / 128.0)):
        path, got = run(fmt, torch.from_numpy(raw).cuda())
        assert path == 1
        conv = pb.normalizeIQ(fmt, raw)          # bit-exact normalizeIQ, then the same front end on CPX
        xd_keep = xd
        xd = torch.from_numpy(conv).cuda()
        _, want = run()
        xd = xd_keep
        for i in range(len(ks)):
            assert rel_rms(got[i], want[i]) <= 1e-12, (fmt, ks[i], rel_rms(got[i], want[i]))


def test_channelizer_1024_channels_bench_shape_vs_oracle():
    """BASELINE config 4 at full width: 1024 channels on the 60 kHz grid, the bench's launch shape (polyphase front end,
    PEBBLE_CHZ_PARTS default, slot grouping by demodulator). 32 sampled channels incl. k = 0, 1, 511, 512, 1022, 1023 are
    compared with the reference chain run once per channel: compiled CDownConvert (MAX_DECSTAGES 12) -> gain restore ->
    compiled CFastFIR -> compiled AGC_OFF -> compiled Demod_AM / Demod_NFM (oracle/_ref). AM channels <= 1e-10; NBFM (float
    PLL state) from sample 0: <= 1e-4 after a lock-time allowance of 1024 output samples (17 ms: while the loop is unlocked
    its phase detector, atan2 of a near-zero sample, amplifies 1e-16 input differences to O(1), so the acquisition path
    itself is not reproducible between any two implementations), and bounded (same order of magnitude) inside it."""
    import torch
    import pebblesdr_b200 as pb
    from oracle import ref, ref_app
    if not (ref_app.dec12_available() and ref_app.available() and ref.available()):
        pytest.skip("oracle/_ref not built")
    nch, n = 1024, 1024 * 2048 * 2
    freqs = [GRID0 + STEP * k for k in range(nch)]
    rng = np.random.default_rng(7)
    sample = sorted(set([0, 1, 511, 512, 1022, 1023] + [int(v) for v in rng.choice(nch, 26, replace=False)]))
    # carriers only where they are checked (+ both neighbours of each, so adjacent-channel leakage is exercised)
    lit = sorted(set(k2 for k in sample for k2 in (k - 1, k, k + 1) if 0 <= k2 < nch))
    t = np.arange(n, dtype=np.float64)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * 0.002
    for k in lit:
        f = freqs[k]
        if k % 2 == 0:
            ph = 2 * np.pi * (f / FS) * t + 5.0 * np.sin(2 * np.pi * (1000.0 / FS) * t)
            x += (1.0 / 64) * np.exp(1j * ph)
        else:
            x += (1.0 / 64) * (1 + 0.5 * np.cos(2 * np.pi * (1000.0 / FS) * t)) * np.exp(2j * np.pi * (f / FS) * t)
    x = np.ascontiguousarray(x)
    cz = pb.Channelizer(nch, FS, BW)
    cz.setFilter(-7500.0, 7500.0)
    for k in range(nch):
        cz.setFrequency(k, freqs[k])
        cz.setDemodMode(k, 2 if k % 2 == 0 else 0)
    nd = n >> 10
    out = torch.zeros(nch, nd, dtype=torch.complex128, device="cuda")
    assert cz.process(torch.from_numpy(x).cuda(), n, out) == nd
    assert cz.frontPath() == 1
    got = out.cpu().numpy()
    gain = 10.0 ** ((10 * 2) / 20.0)
    for k in sample:
        dec = _ref12_channel(x, freqs[k]) * gain                      # receiver.cpp:908, 935-938
        ff = ref.CFastFIR(); ff.SetupParameters(-7500.0, 7500.0, 0.0, 60000.0)
        agc = ref_app.AGC(60000, 2048); agc.setAgcMode(0, 20)
        dem = ref_app.Demod_NFM(60000, 2048) if k % 2 == 0 else ref_app.Demod_AM(60000, 2048)
        want = []
        for i in range(0, nd, 2048):
            bp = np.zeros(2048, dtype=np.complex128)
            assert ff.ProcessData(2048, np.ascontiguousarray(dec[i:i + 2048]), bp) == 2048
            y = agc.processBlock(bp)
            want.append(dem.processBlockNCO(y) if k % 2 == 0 else dem.processBlockFiltered(y))
        want = np.concatenate(want)
        if k % 2 == 0:
            assert rel_rms(got[k][1024:], want[1024:]) <= 1e-4, (k, rel_rms(got[k][1024:], want[1024:]))
            # lock-time allowance: during the first 1024 samples (17 ms) the band-pass is still filling and the loop runs on a
            # vanishing input, where it amplifies 1e-16 differences of the front end (measured 0.15-0.51 relative RMS over
            # this window, varying with the front-end kernel's summation order): an unlocked loop has no sample-level parity.
            # Required there: the same bounded output, and agreement once the carrier is in (last quarter of the window).
            print("nbfm lock window", k, rel_rms(got[k][:1024], want[:1024]), rel_rms(got[k][768:1024], want[768:1024]))
            assert np.abs(got[k][:1024]).max() <= 1.5 * np.abs(want[:1024]).max() + 1e-3, k
            assert rel_rms(got[k][768:1024], want[768:1024]) <= 1e-2, (k, rel_rms(got[k][768:1024], want[768:1024]))
        else:
            assert rel_rms(got[k], want) <= 1e-10, (k, rel_rms(got[k], want))


def test_squelch_is_a_per_channel_enable_and_keeps_demodulator_state():
    """receiver.cpp:955-965: below the squelch level a frame returns BEFORE the AGC and the demodulator — their state does not
    advance, the down-converter and band-pass keep running. Three one-frame calls A, B, C with channels 0 (NBFM) and 1 (AM)
    squelched during B only:
      * open channels are bit-identical to an unsquelched channelizer in all three calls;
      * squelched channels give zeros in B;
      * in C they give what the single-stream Demod operator (pinned to the compiled reference in test_gpu_demod.py) gives
        when it sees the band-pass output of A and then of C — taken from a third channelizer whose channels 0, 1 are
        pass-through (USB), i.e. its rows ARE the band-pass output."""
    import torch
    import pebblesdr_b200 as pb
    nch, frame = 6, 2048
    n1 = frame << 10
    freqs = [-30.69e6 + 60e3 * (k * 170 + 3) for k in range(nch)]
    modes = [2, 0, 2, 0, 7, 0]
    x = torch.from_numpy(_capture(3 * n1, freqs, seed=99)).cuda()

    def make(ms):
        cz = pb.Channelizer(nch, FS, BW)
        cz.setFilter(-7500.0, 7500.0)
        for k in range(nch):
            cz.setFrequency(k, freqs[k]); cz.setDemodMode(k, ms[k])
        return cz

    def run(cz, masks):
        outs = []
        for i, m in enumerate(masks):
            cz.setSquelch(m)
            o = torch.full((nch, frame), 7.0 + 7.0j, dtype=torch.complex128, device="cuda")
            assert cz.process(x[i * n1:(i + 1) * n1], n1, o) == frame
            outs.append(o.cpu().numpy())
        return outs

    closed = np.array([0, 0, 1, 1, 1, 1], dtype=np.uint8)
    g0 = run(make(modes), [None, None, None])
    g1 = run(make(modes), [None, closed, np.ones(nch, dtype=np.uint8)])
    bp = run(make([7, 7] + modes[2:]), [None, None, None])
    for i in range(3):
        assert np.array_equal(g1[i][2:], g0[i][2:])                   # open channels: untouched
    assert np.array_equal(g1[0][:2], g0[0][:2])
    assert np.all(g1[1][:2] == 0)
    assert not np.array_equal(g1[2][:2], g0[2][:2])                   # the gap changed the carried state
    for ch, mode, tol in ((0, pb.Demod.dmFMN, 1e-4), (1, pb.Demod.dmAM, 1e-10)):
        d = pb.Demod(60000, frame)
        d.setDemodMode(mode)
        if mode == pb.Demod.dmAM:
            d.setBandwidth(16000.0)
        a = d.processBlock(np.ascontiguousarray(bp[0][ch]), frame).copy()
        c = d.processBlock(np.ascontiguousarray(bp[2][ch]), frame).copy()
        assert rel_rms(g1[0][ch], a) <= tol, (ch, rel_rms(g1[0][ch], a))
        assert rel_rms(g1[2][ch], c) <= tol, (ch, rel_rms(g1[2][ch], c))


def test_squelch_decision_from_one_wideband_spectrum():
    """pebble_channelizer_squelch_from_spectrum == fdEstimate per channel (pinned to the compiled reference in
    test_gpu_prefilter.py) with the channelizer's own band-pass edges and channel centres; closed channels then give zeros."""
    import torch
    import pebblesdr_b200 as pb
    nch, frame, bins = 8, 2048, 8192
    n1 = frame << 10
    freqs = [-30.69e6 + 60e3 * (k * 120 + 5) for k in range(nch)]
    rng = np.random.default_rng(5)
    db = -110.0 + 3.0 * rng.standard_normal(bins)
    for k in (1, 4, 6):                                               # carriers on three of the channels
        b = int(bins / 2 + freqs[k] / (int(FS) // bins))
        db[b - 1:b + 2] = -40.0
    d_db = torch.from_numpy(np.clip(db, -120.0, 0.0)).cuda()
    cz = pb.Channelizer(nch, FS, BW)
    cz.setFilter(-7500.0, 7500.0)
    for k in range(nch):
        cz.setFrequency(k, freqs[k]); cz.setDemodMode(k, 0)
    want = pb.fdEstimate(d_db, FS, -7500.0, 7500.0, freqs, squelchDb=-80.0)[:, 4]
    assert want.sum() == 3 and cz.squelchFromSpectrum(d_db, FS, -80.0) == 3
    x = torch.from_numpy(_capture(n1, freqs, seed=7)).cuda()
    o = torch.full((nch, frame), 7.0 + 7.0j, dtype=torch.complex128, device="cuda")
    assert cz.process(x, n1, o) == frame
    o = o.cpu().numpy()
    for k in range(nch):
        assert (np.abs(o[k]).max() > 0) == bool(want[k]), k
