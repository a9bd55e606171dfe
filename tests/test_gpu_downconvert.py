"""GPU parity: CDownConvert (fused CUDA kernel, through the C ABI) vs the reference CDownConvert.
Tolerance: relative RMS <= 1e-10; stage plan, output rate and sample counts exact."""
import numpy as np
import pytest

from conftest import rel_rms, iq_noise_tones

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _pair(rate, bw, freq):
    import pebblesdr_b200 as pb
    from oracle import ref
    g, r = pb.CDownConvert(), ref.CDownConvert()
    og, orr = g.SetDataRate(rate, bw), r.SetDataRate(rate, bw)
    assert og == orr
    g.SetFrequency(freq)
    r.SetFrequency(freq)
    return g, r


def _feed(obj, x, sizes, S):
    outs, pos = [], 0
    for n in sizes:
        out = np.zeros(max(n >> S, 1), dtype=np.complex128)
        k = obj.ProcessData(n, np.ascontiguousarray(x[pos:pos + n]), out)
        assert k == n >> S
        outs.append(out[:k].copy())
        pos += n
    return np.concatenate(outs)


@pytest.mark.parametrize("rate,expect", [
    (2e6, [11, 11, 15, 19, 31]),                      # 2 MSPS -> /32, 62 500 Hz
    (10e6, [3, 11, 11, 11, 11, 15, 27]),              # 10 MSPS -> /128, 78 125 Hz
    (61.44e6, [3, 3, 3, 11, 11, 11, 11, 15, 19, 35]),  # 61.44 MSPS -> /1024 (reference overflows MAX_DECSTAGES, F8)
])
def test_stage_plan(rate, expect):
    import pebblesdr_b200 as pb
    g = pb.CDownConvert()
    out = g.SetDataRate(rate, 15000.0)
    assert g.StageLengths() == expect
    assert out == rate / (1 << len(expect))


def test_config2_10msps_chain():
    # BASELINE config 2 front end: 10 MSPS, +1.2 MHz, fed in 32768-sample calls
    n = 32768 * 16
    t = np.arange(n)
    x = iq_noise_tones(n, 2345, 10e6, [(-2e6, 0.5)], noise=0.01)
    x += 0.3 * (1 + 0.5 * np.cos(2 * np.pi * 1e3 / 10e6 * t)) * np.exp(2j * np.pi * 1.2e6 / 10e6 * t)
    g, r = _pair(10e6, 15000.0, 1.2e6)
    S = g.NumStages()
    yg = _feed(g, x, [32768] * 16, S)
    yr = _feed(r, x, [32768] * 16, S)
    assert rel_rms(yg, yr) <= TOL


def test_one_call_equals_many_calls():
    # the fused kernel splits one big call into many CTA segments with pre-roll: must equal 32768-sample calls
    n = 32768 * 32
    x = iq_noise_tones(n, 11, 10e6, [(1.2e6 + 500.0, 0.4), (3e6, 0.4)])
    g1, r = _pair(10e6, 15000.0, 1.2e6)
    g2, _ = _pair(10e6, 15000.0, 1.2e6)
    S = g1.NumStages()
    y1 = _feed(g1, x, [n], S)
    y2 = _feed(g2, x, [32768] * 32, S)
    yr = _feed(r, x, [32768] * 32, S)
    assert rel_rms(y1, yr) <= TOL
    assert rel_rms(y2, yr) <= TOL
    assert rel_rms(y1, y2) <= 1e-12


def test_2msps_and_frequency_change_keeps_phase():
    # SetFrequency does not reset the oscillator phase (downconvert.cpp:102-113)
    n = 8192 * 8
    x = iq_noise_tones(n, 21, 2e6, [(3e5, 0.5), (-4e5, 0.3)])
    g, r = _pair(2e6, 15000.0, 3e5)
    S = g.NumStages()
    ya = _feed(g, x[: n // 2], [8192] * 4, S)
    ra = _feed(r, x[: n // 2], [8192] * 4, S)
    g.SetFrequency(-4e5)
    r.SetFrequency(-4e5)
    yb = _feed(g, x[n // 2:], [8192] * 4, S)
    rb = _feed(r, x[n // 2:], [8192] * 4, S)
    assert rel_rms(np.concatenate([ya, yb]), np.concatenate([ra, rb])) <= TOL


def test_input_not_clobbered_and_bad_length():
    import pebblesdr_b200 as pb
    g, _ = _pair(10e6, 15000.0, 1.2e6)
    x = iq_noise_tones(4096, 1, 10e6, [(1.2e6, 0.5)])
    keep = x.copy()
    out = np.zeros(64, complex)
    g.ProcessData(4096, x, out)
    assert np.array_equal(x, keep)          # documented deviation F6: input left untouched
    with pytest.raises(pb.PebbleError):
        g.ProcessData(100, x, out)          # not a multiple of 2^7


def test_device_resident_large_matches_chunked_reference():
    # 2^24 samples resident on the device in ONE call vs the reference fed 32768 at a time
    import torch
    n = 1 << 24
    x = iq_noise_tones(n, 77, 10e6, [(1.2e6 + 800.0, 0.3), (-2e6, 0.5)], noise=0.01)
    g, r = _pair(10e6, 15000.0, 1.2e6)
    S = g.NumStages()
    xd = torch.from_numpy(x).cuda()
    yd = torch.zeros(n >> S, dtype=torch.complex128, device="cuda")
    assert g.ProcessData(n, xd, yd) == n >> S
    yg = yd.cpu().numpy()
    yr = _feed(r, x, [32768] * (n // 32768), S)
    assert rel_rms(yg, yr) <= TOL


def test_reference_call_shape_2048_and_ragged_small_calls_10msps():
    """The reference's real call shape is framesPerBuffer = 2048 samples per processIQData (16 outputs per call), and the
    minimal legal call is 2^S = 128 samples: both hit dc_pipe_kernel's single-partial-chunk path and the state hand-over."""
    sizes = [2048] * 24 + [128, 384, 2048, 128, 128, 1024, 4096 + 128]
    sizes.append(3 * 32768 - sum(sizes))
    n = sum(sizes)
    assert n % 32768 == 0 and sizes[-1] % 128 == 0
    x = iq_noise_tones(n, 31, 10e6, [(1.2e6 + 700.0, 0.4), (-2e6, 0.4)], noise=0.02)
    g, r = _pair(10e6, 15000.0, 1.2e6)
    S = g.NumStages()
    yg = _feed(g, x, sizes, S)
    yr = _feed(r, x, [32768] * (n // 32768), S)   # the reference filters every stage only for calls this long (see below)
    assert rel_rms(yg, yr) <= TOL


def test_61msps_10_stage_chain_vs_reference_max_decstages_12():
    """BASELINE config 4's chain (CIC3 x3, HB11 x4, HB15, HB19, HB35) against the reference compiled with MAX_DECSTAGES 12
    (downconvert.h:23; oracle/build_ref.sh step 6), incl. 2048-sample calls on the GPU side."""
    import pebblesdr_b200 as pb
    from oracle import ref_app
    if not ref_app.dec12_available():
        pytest.skip("oracle/_ref/libpebbleref_dec12.so not built")
    n = 65536 * 4
    f = -30.69e6 + 700 * 60e3
    x = iq_noise_tones(n, 32, 61.44e6, [(f + 1500.0, 0.3), (f + 60e3, 0.3), (9e6, 0.4)], noise=0.02)
    for sizes in ([32768] * 8, [2048] * 16 + [1024, 3072, 4096] + [32768 - 8192] + [32768] * 5 + [1024] * 32):
        g, r = pb.CDownConvert(), ref_app.CDownConvert12()
        assert g.SetDataRate(61.44e6, 15000.0) == r.SetDataRate(61.44e6, 15000.0) == 60000.0
        g.SetFrequency(f); r.SetFrequency(f)
        assert sum(sizes) == n
        yg = _feed(g, x, sizes, 10)
        # 65536-sample reference calls: the 35-tap stage then receives 128 >= 2 * 34 samples (see the short-call test below)
        yr = _feed(r, x, [65536] * 4, 10)
        assert rel_rms(yg, yr) <= TOL


@pytest.mark.parametrize("rate,bw,simple,expect", [
    (2e6, 15000.0, True, [51, 51, 51]),            # SetDataRateSimple: 51-tap stages only (downconvert.cpp:220-244)
    (48000.0, 2000.0, False, [15, 23]),
    (48000.0, 6500.0, False, [39]),
    (3.2e6, 30000.0, False, [11, 11, 15, 23, 43]),
    (96000.0, 15000.0, False, [47]),
    (48000.0, 8000.0, False, [51]),
    (250000.0, 30000.0, False, [31]),
])
def test_every_generic_halfband_length_runs_on_the_gpu(rate, bw, simple, expect):
    """each of HB15...HB51 (filtercoef.h:52-426, F7's doubled tap 0 included) executed on the device."""
    import pebblesdr_b200 as pb
    from oracle import ref
    g, r = pb.CDownConvert(), ref.CDownConvert()
    if simple:
        assert g.SetDataRateSimple(rate, bw) == r.SetDataRateSimple(rate, bw)
    else:
        assert g.SetDataRate(rate, bw) == r.SetDataRate(rate, bw)
    assert g.StageLengths() == expect
    f = rate * 0.11
    g.SetFrequency(f); r.SetFrequency(f)
    S = len(expect)
    n = 4096 * 6
    x = iq_noise_tones(n, 33, rate, [(f + 0.01 * rate / (1 << S), 0.4), (-0.3 * rate, 0.4)], noise=0.05)
    yg = _feed(g, x, [4096, 8192, 512, 3584, 8192], S)
    yr = _feed(r, x, [4096] * 6, S)
    assert rel_rms(yg, yr) <= TOL


def test_two_handles_sharing_a_kernel_do_not_lower_each_others_shared_memory_limit():
    """The dynamic shared-memory limit is an attribute of the kernel, not of a handle: a 3 x 51-tap chain (large stage arrays)
    and a one-stage chain run the same generic kernel with different sizes, interleaved. Regression for a launch failure
    ("invalid argument") once the smaller handle had set the limit after the larger one."""
    import pebblesdr_b200 as pb
    from oracle import ref
    big, small = pb.CDownConvert(), pb.CDownConvert()
    rb, rs = ref.CDownConvert(), ref.CDownConvert()
    assert big.SetDataRateSimple(2e6, 15000.0) == rb.SetDataRateSimple(2e6, 15000.0)
    assert small.SetDataRate(250000.0, 30000.0) == rs.SetDataRate(250000.0, 30000.0)
    for o in (big, rb):
        o.SetFrequency(2e5)
    for o in (small, rs):
        o.SetFrequency(3e4)
    xb = iq_noise_tones(4096 * 6, 35, 2e6, [(2.01e5, 0.4)], noise=0.05)
    xs = iq_noise_tones(4096 * 6, 36, 250000.0, [(3.05e4, 0.4)], noise=0.05)
    yb, ys = [], []
    for i in range(0, 4096 * 6, 4096):                       # big, small, big, small, ...
        yb.append(_feed(big, xb[i:i + 4096], [4096], 3)); ys.append(_feed(small, xs[i:i + 4096], [4096], 1))
    assert rel_rms(np.concatenate(yb), _feed(rb, xb, [4096] * 6, 3)) <= TOL
    assert rel_rms(np.concatenate(ys), _feed(rs, xs, [4096] * 6, 1)) <= TOL


def test_short_call_rule_documented_deviation_and_strict_mode():
    """downconvert.cpp:362-363: a generic half-band stage that receives fewer samples than it has taps returns n/2 WITHOUT
    filtering (the caller's buffer is left as it is, so the 'output' is the stage's leading input samples); and with fewer
    than 2*(taps-1) samples it re-reads its carried history from the caller's buffer AFTER the in-place outputs overwrote
    it (:387-390; CDownConvert always calls DecBy2 in place, :325). So the reference's output depends on how the stream is
    cut into calls (first two assertions). This library filters every stage for every call length: short calls equal long
    calls of the same stream. pebble_downconvert_set_strict_lengths(h, 1) turns such calls into PEBBLE_ERR_ARG for
    callers that want the reference's precondition enforced."""
    import pebblesdr_b200 as pb
    from oracle import ref
    n = 32768
    x = iq_noise_tones(n, 34, 10e6, [(1.2e6 + 700.0, 0.4)], noise=0.02)
    r_long, r_short = ref.CDownConvert(), ref.CDownConvert()
    for r in (r_long, r_short):
        r.SetDataRate(10e6, 15000.0); r.SetFrequency(1.2e6)
    y_long = _feed(r_long, x, [n], 7)
    y_short = _feed(r_short, x, [1024] * 32, 7)          # stage 6 (27 taps) sees 16 samples per call: not filtered
    assert rel_rms(y_short, y_long) > 1e-3
    r_mid = ref.CDownConvert()
    r_mid.SetDataRate(10e6, 15000.0); r_mid.SetFrequency(1.2e6)
    y_mid = _feed(r_mid, x, [2048] * 16, 7)               # 32 samples into the 27-tap stage: filtered, but 32 < 2 * 26
    assert 1e-9 < rel_rms(y_mid, y_long) < 1e-2
    g, _ = _pair(10e6, 15000.0, 1.2e6)
    assert rel_rms(_feed(g, x, [1024] * 32, 7), y_long) <= TOL
    g2, _ = _pair(10e6, 15000.0, 1.2e6)
    pb.lib.pebble_downconvert_set_strict_lengths(g2._h, 1)
    out = np.zeros(256, complex)
    with pytest.raises(pb.PebbleError):
        g2.ProcessData(1024, np.ascontiguousarray(x[:1024]), out)
    with pytest.raises(pb.PebbleError):
        g2.ProcessData(2048, np.ascontiguousarray(x[:2048]), out)            # 32 < 2 * 26
    assert g2.ProcessData(4096, np.ascontiguousarray(x[:4096]), out) == 32   # 4096 >> 6 = 64 >= 52: fine
