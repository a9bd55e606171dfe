"""GPU parity: device sample formats. DeviceInterfaceBase::normalizeIQ (pebblelib/deviceinterfacebase.cpp:648-878)
runs on the host in the reference and hands CPX to Receiver::processIQData; here the raw int16 / uint8 / int8 / float samples
are what the chain takes, converted in registers inside the mixer's load stage.

  * the stand-alone conversion (pebble_normalize_iq) is BIT-IDENTICAL to the compiled reference for every code, all four IQ
    orders, both `reverse` values and non-trivial gains;
  * CDownConvert and the receiver chain fed raw samples equal "reference normalizeIQ -> reference CDownConvert (-> oracle
    chain)" within the chain tolerance 1e-10 relative RMS (the scale rides on the oscillator amplitude: one rounding apart
    unless it is a power of two);
  * orders/chains with no fused kernel take the convert-then-mix path and must agree as well."""
import numpy as np
import pytest

from conftest import rel_rms

pytestmark = pytest.mark.gpu
TOL = 1e-10
FMT_NP = {1: np.int16, 2: np.uint8, 3: np.int8, 4: np.float32}


def _raw(fmt, n, seed, fs, tones, noise=0.02):
    """Interleaved raw samples of a synthetic IQ stream quantised to the format's codes."""
    rng = np.random.default_rng(seed)
    t = np.arange(n)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * noise
    for f, a in tones:
        x = x + a * np.exp(2j * np.pi * (f / fs) * t)
    f2 = np.stack([x.real, x.imag], axis=1).reshape(-1)
    if fmt == 1:
        return np.clip(np.round(f2 * 32767), -32768, 32767).astype(np.int16)
    if fmt == 2:
        return (np.clip(np.round(f2 * 127), -128, 127) + 128).astype(np.uint8)
    if fmt == 3:
        return np.clip(np.round(f2 * 127), -128, 127).astype(np.int8)
    if fmt == 4:
        return f2.astype(np.float32)
    return np.ascontiguousarray(x)


def _ref_fmt(fmt):
    from oracle import ref_app
    return {1: ref_app.FMT_CPX16, 2: ref_app.FMT_CPXU8, 3: ref_app.FMT_CPX8, 4: ref_app.FMT_CPXFLOAT, 0: ref_app.FMT_CPX}[fmt]


def test_normalize_iq_bit_exact_every_code():
    import pebblesdr_b200 as pb
    from oracle import ref_app
    i8 = np.arange(-128, 128, dtype=np.int8)
    u8 = np.arange(0, 256, dtype=np.uint8)
    i16 = np.arange(-32768, 32768, dtype=np.int32).astype(np.int16)
    f32 = np.concatenate([np.linspace(-1.5, 1.5, 8191, dtype=np.float32), np.array([0.0, -0.0, 1e-40, 3.4e38], dtype=np.float32)])
    raws = {3: np.stack([i8, i8[::-1]], axis=1).reshape(-1), 2: np.stack([u8, u8[::-1]], axis=1).reshape(-1),
            1: np.stack([i16, i16[::-1]], axis=1).reshape(-1), 4: np.stack([f32, f32[::-1]], axis=1).reshape(-1)}
    raws[0] = np.ascontiguousarray(raws[4][0::2].astype(np.float64) + 1j * raws[4][1::2].astype(np.float64))
    for fmt, raw in raws.items():
        for order in range(4):
            for rev in (False, True):
                for ug, ng in ((1.0, 1.0), (0.75, 1.0 / 3.0)):
                    got = pb.normalizeIQ(fmt, raw, order, ug, ng, rev)
                    rraw = raw if fmt else np.stack([raw.real, raw.imag], axis=1).reshape(-1)
                    want = ref_app.normalizeIQ(_ref_fmt(fmt), rraw, order, ug, ng, rev)
                    assert np.array_equal(got, want), (fmt, order, rev, ug)


def test_normalize_iq_device_tensor_matches_host_path():
    import torch
    import pebblesdr_b200 as pb
    raw = _raw(1, 100003, 5, 1e6, [(1e5, 0.4)])
    a = pb.normalizeIQ(pb.IQ_CS16, raw)
    b = pb.normalizeIQ(pb.IQ_CS16, torch.from_numpy(raw).cuda()).cpu().numpy()
    assert np.array_equal(a, b)


@pytest.mark.parametrize("fmt", [1, 2, 3, 4])
@pytest.mark.parametrize("gains", [(1.0, 1.0), (0.6, 1.7)])
def test_downconvert_raw_formats_10msps_fused(fmt, gains):
    """10 MSPS chain (dc_pipe_kernel<..., FMT>): raw samples in, vs reference normalizeIQ + reference CDownConvert."""
    import pebblesdr_b200 as pb
    from oracle import ref, ref_app
    n = 32768 * 12
    raw = _raw(fmt, n, 40 + fmt, 10e6, [(1.2015e6, 0.3), (-2e6, 0.5)])
    x = ref_app.normalizeIQ(_ref_fmt(fmt), raw, 0, gains[0], gains[1])
    g, r = pb.CDownConvert(), ref.CDownConvert()
    assert g.SetDataRate(10e6, 15000.0) == r.SetDataRate(10e6, 15000.0)
    g.SetFrequency(1.2e6); r.SetFrequency(1.2e6)
    g.SetInputFormat(fmt, pb.IQO_IQ, gains[0], gains[1])
    S = g.NumStages()
    yg, pos = [], 0
    for m in [32768] * 4 + [128, 384, 2048, 32768 - 2560] + [32768 * 7]:   # small and ragged calls included
        a = np.zeros(m >> S, dtype=np.complex128)
        assert g.ProcessData(m, np.ascontiguousarray(raw[2 * pos:2 * (pos + m)]), a) == m >> S
        yg.append(a); pos += m
    assert pos == n
    yr = []
    for q in range(0, n, 32768):   # the reference only filters every stage for calls this long (downconvert.cpp:362-363)
        b = np.zeros(32768 >> S, dtype=np.complex128)
        assert r.ProcessData(32768, x[q:q + 32768], b) == 32768 >> S
        yr.append(b)
    assert rel_rms(np.concatenate(yg), np.concatenate(yr)) <= TOL


@pytest.mark.parametrize("fmt,order", [(1, 1), (2, 2), (3, 3), (0, 1)])
def test_downconvert_raw_formats_unfused_orders(fmt, order):
    """QI / I-only / Q-only orders (and CPX with a swap) have no fused kernel: convert-then-mix, same answer."""
    import pebblesdr_b200 as pb
    from oracle import ref, ref_app
    n = 32768 * 4
    raw = _raw(fmt, n, 60 + fmt, 10e6, [(1.2015e6, 0.3), (-2e6, 0.5)])
    rraw = raw if fmt else np.stack([raw.real, raw.imag], axis=1).reshape(-1)
    x = ref_app.normalizeIQ(_ref_fmt(fmt), rraw, order, 1.0, 1.0)
    g, r = pb.CDownConvert(), ref.CDownConvert()
    g.SetDataRate(10e6, 15000.0); r.SetDataRate(10e6, 15000.0)
    g.SetFrequency(1.2e6); r.SetFrequency(1.2e6)
    g.SetInputFormat(fmt, order)
    S = g.NumStages()
    a = np.zeros(n >> S, dtype=np.complex128)
    assert g.ProcessData(n, raw, a) == n >> S
    b = np.zeros(n >> S, dtype=np.complex128)
    for q in range(0, n, 32768):
        r.ProcessData(32768, x[q:q + 32768], b[q >> S:])
    assert rel_rms(a, b) <= TOL


def test_downconvert_raw_format_other_chain_2msps():
    """A chain without a fused-format kernel (2 MSPS: HB11,HB11,HB15,HB19,HB31) converts into scratch first."""
    import pebblesdr_b200 as pb
    from oracle import ref, ref_app
    n = 8192 * 6
    raw = _raw(1, n, 77, 2e6, [(-3.49e5, 0.4), (6e5, 0.5)])
    x = ref_app.normalizeIQ(ref_app.FMT_CPX16, raw)
    g, r = pb.CDownConvert(), ref.CDownConvert()
    g.SetDataRate(2e6, 15000.0); r.SetDataRate(2e6, 15000.0)
    g.SetFrequency(-3.5e5); r.SetFrequency(-3.5e5)
    g.SetInputFormat(pb.IQ_CS16)
    S = g.NumStages()
    a, b = np.zeros(n >> S, dtype=np.complex128), np.zeros(n >> S, dtype=np.complex128)
    assert g.ProcessData(n, raw, a) == n >> S
    for q in range(0, n, 8192):
        r.ProcessData(8192, x[q:q + 8192], b[q >> S:])
    assert rel_rms(a, b) <= TOL


@pytest.mark.parametrize("fmt", [1, 2])
def test_receiver_chain_raw_samples_vs_oracle_chain(fmt):
    """BASELINE config 2 chain fed int16 / uint8 IQ == reference normalizeIQ -> oracle chain on the converted stream."""
    import torch
    import pebblesdr_b200 as pb
    from oracle import port, ref_app
    n = 32768 * 20
    t = np.arange(n)
    rng = np.random.default_rng(99)
    x = 0.3 * (1 + 0.5 * np.cos(2 * np.pi * 1e3 / 10e6 * t)) * np.exp(2j * np.pi * 1.2015e6 / 10e6 * t)
    x = x + 0.5 * np.exp(-2j * np.pi * 2e6 / 10e6 * t) + (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * 0.01
    f2 = np.stack([x.real, x.imag], axis=1).reshape(-1)
    raw = np.round(f2 * 20000).astype(np.int16) if fmt == 1 else (np.round(f2 * 80) + 128).astype(np.uint8)
    xc = ref_app.normalizeIQ(_ref_fmt(fmt), raw)
    rx = pb.Receiver(10e6, 2048)
    assert rx.setDataRate(15000.0) == 78125.0
    rx.setMixer(1.2e6); rx.setFilter(-5000.0, 5000.0); rx.setDemodMode(pb.Demod.dmAM); rx.setAgc(0, 20)
    rx.setDemodBandwidth(10000.0)
    rx.setInputFormat(fmt)
    k, y = rx.processIQData(torch.from_numpy(raw).cuda(), n)
    y = y[:k].cpu().numpy()
    o = port.Receiver(10e6, 2048, 15000.0)
    o.config(1.2e6, -5000.0, 5000.0, 0, 0, 20)
    o.setAmBandwidth(10000.0)
    yo = np.concatenate([o.processIQData(xc[i:i + 32768], 32768) for i in range(0, n, 32768)])
    assert k == len(yo) > 0
    assert rel_rms(y, yo) <= TOL
    # host-pointer call shape (the reference's 2048-sample calls), raw samples from a numpy array
    rx2 = pb.Receiver(10e6, 2048)
    rx2.setDataRate(15000.0); rx2.setMixer(1.2e6); rx2.setFilter(-5000.0, 5000.0); rx2.setDemodMode(pb.Demod.dmAM)
    rx2.setAgc(0, 20); rx2.setDemodBandwidth(10000.0); rx2.setInputFormat(fmt)
    outs = []
    for i in range(0, 32768 * 10, 2048):
        kk, yy = rx2.processIQData(np.ascontiguousarray(raw[2 * i:2 * (i + 2048)]), 2048)
        outs.append(yy[:kk].copy())
    y2 = np.concatenate(outs)
    assert rel_rms(y2, yo[:len(y2)]) <= TOL
