"""CPU test (no GPU): the host planner of the fractional resampler. CFractResampler::Resample's output times are a chain of
rounded double additions (m_FloatTime += dt, pebblelib/fractresampler.cpp:174); the library walks them in O(binades) steps
(csrc/resampler.cu rs_segments) and must reproduce the chain bit for bit: output counts, every output time and the carried
m_FloatTime, against (a) the literal loop and (b) the output counts of the COMPILED reference (oracle/_ref)."""
import numpy as np
import pytest

from oracle import ref_app


def _loop(t, dt, n):
    out = []
    while int(t) < n:          # fractresampler.cpp:159
        out.append(t)
        t = t + dt             # :174
    return out, t - n          # :177


@pytest.mark.parametrize("dt", [78125.0 / 48000.0, 60000.0 / 48000.0, 62500.0 / 44100.0, 0.5, 2.0, 1.0000001, 0.9999999, 3.7, 0.31])
def test_planner_equals_literal_loop(dt):
    import pebblesdr_b200 as pb
    rng = np.random.default_rng(int(dt * 1000))
    for trial in range(150):
        n = int(rng.integers(1, 6000))
        t0 = 0.0 if trial % 4 == 0 else float(rng.uniform(0, max(dt, 1.0)))
        want, te = _loop(t0, dt, n)
        cnt, te2, times = pb.resampler_plan(t0, dt, n)
        assert cnt == len(want) and te2 == te
        assert np.array_equal(times, np.array(want))


def test_planner_chained_frames_and_ragged_calls():
    import pebblesdr_b200 as pb
    for dt, sizes in ((78125.0 / 48000.0, [2048] * 400), (60000.0 / 48000.0, [2048, 1, 17, 4095, 2048, 333] * 40)):
        t = t2 = 0.0
        for n in sizes:
            w, t = _loop(t, dt, n)
            c, t2, _ = pb.resampler_plan(t2, dt, n)
            assert c == len(w) and t == t2


@pytest.mark.skipif(not ref_app.available(), reason="oracle/_ref/libpebbleref_app.so not built (needs /root/reference)")
@pytest.mark.parametrize("rate,sizes", [(78125.0 / 48000.0, [2048] * 12), (60000.0 / 48000.0, [2048, 100, 1, 3000, 29, 2048])])
def test_planner_counts_equal_compiled_reference(rate, sizes):
    import pebblesdr_b200 as pb
    r = ref_app.CFractResampler(4096)
    rng = np.random.default_rng(5)
    t = 0.0
    for n in sizes:
        x = rng.standard_normal(n) + 1j * rng.standard_normal(n)
        want = len(r.Resample(x, rate))
        cnt, t, _ = pb.resampler_plan(t, rate, n)
        assert cnt == want
