"""GPU parity for the small SURVEY rows: a1 CpxUtil vector helpers (pebblelib/cpx.cpp:35-112) and a17 NCO
(pebblelib/nco.cpp:16-81), both against the COMPILED reference (oracle/_ref/libpebbleref.so)."""
import numpy as np
import pytest

from conftest import rel_rms, iq_noise_tones

pytestmark = pytest.mark.gpu


def test_cpxutil_device_helpers_bit_exact():
    import torch
    import pebblesdr_b200 as pb
    from oracle import ref
    n = 100003
    a = iq_noise_tones(n, 1, 1e6, [(1e5, 0.7)], noise=0.3)
    b = iq_noise_tones(n, 2, 1e6, [(-2e5, 0.4)], noise=0.5)
    da, db = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
    for name, kw in (("copy", {}), ("scale", {"s": 0.123456789}), ("mult", {"b": 1}), ("add", {"b": 1}), ("mag", {}), ("sqrmag", {}),
                     ("clear", {}), ("decimate", {"s": 2}), ("decimate", {"s": 7})):
        bb = b if "b" in kw else None
        want = ref.cpx_op(name, a, bb, kw.get("s", 0.0))
        got = pb.cpx_util(name, da, db if "b" in kw else None, kw.get("s", 0.0)).cpu().numpy()
        assert got.shape == want.shape and np.array_equal(got, want), name


@pytest.mark.parametrize("fs,f", [(48000, 1000.0), (48000, -30000.0), (48000, 23999.0), (48000, 24000.0), (2000000, 1.5e6), (96000, -47999.5)])
def test_nco_nyquist_clamp_and_gen_single(fs, f):
    """NCO::setFrequency limits |f| to (fs / 2) - 1 (integer arithmetic, nco.cpp:21-26); genSingle = gain * nextSample()."""
    import pebblesdr_b200 as pb
    from oracle import ref
    g, r = pb.NCO(fs, 2048), ref.NCO(fs, 2048)
    got_f = g.setFrequency(f)
    r.setFrequency(f)
    ny = fs // 2 - 1
    assert got_f == max(-ny, min(ny, f)) == g.frequency()
    n = 2048 * 6
    base = iq_noise_tones(n, 3, fs, [], noise=0.1)
    for mix in (False, True):
        a, b = base.copy(), base.copy()
        for i in range(0, n, 2048):           # consecutive calls continue the oscillator
            r.genSingle(a[i:], 2048, 0.5, mix)
            assert g.genSingle(b[i:i + 2048], 2048, 0.5, mix) == 2048
        assert rel_rms(b, a) <= 1e-12
    g.setFrequency(f)                          # restart at (1, 0) (nco.cpp:40-41)
    r.setFrequency(f)
    a, b = np.zeros(64, complex), np.zeros(64, complex)
    r.genSingle(a, 64, 1.0, False); g.genSingle(b, 64, 1.0, False)
    assert rel_rms(b, a) <= 1e-13
