"""CPU tests (no GPU): pin the application-side part of the oracle to the COMPILED reference.

oracle/_ref/libpebbleref_app.so is application/{agc,demod/demod_am,demod/demod_nfm,...}.cpp compiled unmodified
against empty Qt stand-ins (oracle/build_ref.sh step 5); libpebbleref_dec12.so is CDownConvert with MAX_DECSTAGES 12
(step 6: the 10-stage 61.44 MSPS chain of BASELINE config 4).

1. tests/golden/app_*.npz (generated from those libraries by tests/golden/make_golden_app.py) must be reproduced by the C
   restatement oracle/pebble_oracle.c — runs everywhere.
2. Where the libraries exist they must reproduce the fixtures too, and the restatement is compared with them live on long
   streams (AGC_LONG's 2000 ms decay needs > 150 k samples at 78 125 S/s to show, agc.cpp:52-82).

The restatement follows the reference statement by statement with the same operation order, so AGC / AM / PLL / FM2 /
normalizeIQ are required to be BIT-IDENTICAL (np.array_equal); CDownConvert within 1e-12 relative RMS."""
import os
import sys

import numpy as np
import pytest

from conftest import rel_rms

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden_app as mga  # noqa: E402
from oracle import port, ref_app  # noqa: E402

needs_ref = pytest.mark.skipif(not (ref_app.available() and ref_app.dec12_available()),
                               reason="oracle/_ref/libpebbleref_app.so not built (needs /root/reference)")
FS, FRAME = mga.FS, mga.FRAME


def _check(name, got, want):
    assert sorted(got.keys()) == sorted(want.files), name
    for key in want.files:
        g, w = np.asarray(got[key]), want[key]
        assert g.shape == w.shape, (name, key)
        if name == "app_dc_61msps" and key == "out":
            assert rel_rms(g, w) <= 1e-12, (name, rel_rms(g, w))
        else:
            assert np.array_equal(g, w), (name, key, rel_rms(g, w))


@pytest.mark.parametrize("name", sorted(mga.APP_CASES))
def test_c_oracle_reproduces_app_golden(name):
    want = np.load(os.path.join(HERE, "golden", name + ".npz"))
    _check(name, mga.run_case(mga.APP_CASES[name], mga.PortImpl()), want)


@needs_ref
@pytest.mark.parametrize("name", sorted(mga.APP_CASES))
def test_reference_app_library_reproduces_golden(name):
    want = np.load(os.path.join(HERE, "golden", name + ".npz"))
    _check(name, mga.run_case(mga.APP_CASES[name], mga.RefImpl()), want)


@needs_ref
@pytest.mark.parametrize("mode,thr,thr2", [(0, 20, 33), (1, 20, 45), (2, 30, 5), (3, 40, 60), (4, 25, 70)])
def test_c_oracle_agc_vs_reference_long_stream(mode, thr, thr2):
    """All five modes over 2^18 samples (3.4 s at 78 125 S/s) with a mid-stream setAgcThreshold."""
    n = 1 << 18
    x = mga.am_signal(n, 100 + mode)
    a = ref_app.AGC(FS, FRAME)
    a.setAgcMode(mode, thr)
    o = port.AGC(FS)
    o.setAgcMode(mode, thr)
    yr, yo = [], []
    for i in range(0, n, FRAME):
        if i == n // 2:
            a.setAgcThreshold(thr2)
            o.setAgcThreshold(thr2)
        yr.append(a.processBlock(x[i:i + FRAME]))
        yo.append(o.processBlock(np.ascontiguousarray(x[i:i + FRAME]), FRAME))
    yr, yo = np.concatenate(yr), np.concatenate(yo)
    assert np.array_equal(yo, yr), rel_rms(yo, yr)
    assert np.abs(yr).max() > 0


@needs_ref
@pytest.mark.parametrize("carrier,dev", [(0.0, 5000.0), (19000.0, 3000.0), (-21000.0, 2000.0), (3000.0, 14000.0)])
def test_c_oracle_nfm_pll_vs_reference_incl_clamp(carrier, dev):
    """processBlockNCO (demod_nfm.cpp:225-257); carriers beyond +-15 kHz pin the NCO frequency to the clamp (:242-245)."""
    n = FRAME * 24
    x = mga.fm_signal(n, 200, carrier=carrier, dev=dev)
    d, o = ref_app.Demod_NFM(FS, FRAME), port.Demod(FS)
    yr = np.concatenate([d.processBlockNCO(x[i:i + FRAME]) for i in range(0, n, FRAME)])
    yo = np.concatenate([o.nfm_pll(np.ascontiguousarray(x[i:i + FRAME]), FRAME) for i in range(0, n, FRAME)])
    assert np.array_equal(yo, yr), rel_rms(yo, yr)
    if abs(carrier) > 15000:
        # locked out: the loop sits on the clamp, so the DC-tracked output decays to ~0 away from the start
        assert np.abs(yr[n // 2:].real).max() < 0.2 * np.abs(yr[:FRAME].real).max()


@needs_ref
def test_c_oracle_am_and_fm2_vs_reference():
    n = FRAME * 16
    x = mga.am_signal(n, 300, steps=False)
    d, o = ref_app.Demod_AM(FS, FRAME), port.Demod(FS)
    for bw in (16000.0, 10000.0, 5000.0):
        d.setBandwidth(bw)
        o.setBandwidth(bw)
        yr = np.concatenate([d.processBlockFiltered(x[i:i + FRAME]) for i in range(0, n, FRAME)])
        yo = np.concatenate([o.am(np.ascontiguousarray(x[i:i + FRAME]), FRAME) for i in range(0, n, FRAME)])
        assert np.array_equal(yo, yr), (bw, rel_rms(yo, yr))
    xf = mga.fm_signal(n, 301)
    d2, o2 = ref_app.Demod_NFM(FS, FRAME), port.Demod(FS)
    d2.processBlockFM2(np.zeros(1, dtype=np.complex128))   # flush the function-static last sample (demod_nfm.cpp:129)
    yr = np.concatenate([d2.processBlockFM2(xf[i:i + FRAME]) for i in range(0, n, FRAME)])
    yo = np.concatenate([o2.fm2(np.ascontiguousarray(xf[i:i + FRAME]), FRAME) for i in range(0, n, FRAME)])
    assert np.array_equal(yo, yr)


@needs_ref
def test_c_oracle_10_stage_chain_vs_reference_max_decstages_12():
    """BASELINE config 4: 61.44 MSPS / 15 kHz plans CIC3 x3, HB11 x4, HB15, HB19, HB35 = 10 stages."""
    from signals import iq
    rate, bw = 61.44e6, 15000.0
    for k in (0, 1, 511, 1023):
        f = -30.69e6 + k * 60e3
        x = iq(32768 * 6, 400 + k, rate, [(f + 2000.0, 0.3), (f - 45e3, 0.2), (7e6, 0.5)], 0.01)
        r, o = ref_app.CDownConvert12(), port.CDownConvert()
        assert r.SetDataRate(rate, bw) == o.SetDataRate(rate, bw) == 60000.0
        assert o.NumStages() == 10 and o.StageLengths() == [3, 3, 3, 11, 11, 11, 11, 15, 19, 35]
        r.SetFrequency(f)
        o.SetFrequency(f)
        yr, yo = [], []
        for i in range(0, len(x), 32768):
            a, b = np.zeros(32, dtype=np.complex128), np.zeros(32, dtype=np.complex128)
            assert r.ProcessData(32768, x[i:i + 32768], a) == o.ProcessData(32768, x[i:i + 32768], b) == 32
            yr.append(a); yo.append(b)
        assert rel_rms(np.concatenate(yo), np.concatenate(yr)) <= 1e-12


@needs_ref
def test_c_oracle_normalize_iq_vs_reference_all_codes():
    """Every int8 / uint8 code and a sweep of int16 codes, all four IQ orders, both `reverse` values: bit-identical."""
    i8 = np.arange(-128, 128, dtype=np.int8)
    raw8 = np.stack([i8, i8[::-1]], axis=1).reshape(-1)
    u8 = np.arange(0, 256, dtype=np.uint8)
    rawu8 = np.stack([u8, u8[::-1]], axis=1).reshape(-1)
    i16 = np.arange(-32768, 32768, 7, dtype=np.int32).astype(np.int16)
    raw16 = np.stack([i16, i16[::-1]], axis=1).reshape(-1)
    f32 = np.linspace(-1.5, 1.5, 4001, dtype=np.float32)
    rawf = np.stack([f32, f32[::-1]], axis=1).reshape(-1)
    for fmt, raw in ((0, raw8), (1, rawu8), (2, raw16), (3, rawf), (4, rawf.astype(np.float64))):
        for order in range(4):
            for rev in (False, True):
                for ug, ng in ((1.0, 1.0), (0.75, 1.0 / 3.0), (2.5, 0.004)):
                    a = ref_app.normalizeIQ(fmt, raw, order, ug, ng, rev)
                    b = port.normalizeIQ(fmt, raw, order, ug, ng, rev)
                    assert np.array_equal(a, b), (fmt, order, rev, ug, ng)
