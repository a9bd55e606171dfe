"""GPU parity: CFastFIR (CUDA, through the C ABI) vs the reference pebblelib/FFTW CFastFIR.
Tolerance: relative RMS <= 1e-10; sample counts and block boundaries exact."""
import numpy as np
import pytest

from conftest import rel_rms, iq_noise_tones

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _pair(lo=-3000.0, hi=3000.0, off=0.0, fs=2e6):
    import pebblesdr_b200 as pb
    from oracle import ref
    g, r = pb.CFastFIR(), ref.CFastFIR()
    assert g.SetupParameters(lo, hi, off, fs) == 0
    r.SetupParameters(lo, hi, off, fs)
    return g, r


def _feed(obj, x, sizes):
    outs, counts, pos = [], [], 0
    for n in sizes:
        out = np.zeros(max(n + 2048, 1), dtype=np.complex128)
        k = obj.ProcessData(n, np.ascontiguousarray(x[pos:pos + n]), out)
        counts.append(k)
        outs.append(out[:k].copy())
        pos += n
    return np.concatenate(outs) if outs else np.zeros(0, complex), counts


def test_config1_bandpass_2msps():
    # BASELINE config 1: 2048-pt overlap-save, +-3 kHz, 2 MSPS, noise + tones at +1 kHz and +50 kHz
    n = 1 << 18
    x = iq_noise_tones(n, 1234, 2e6, [(1e3, 0.5), (50e3, 0.5)])
    g, r = _pair()
    yg, cg = _feed(g, x, [n])
    yr, cr = _feed(r, x, [n])
    assert cg == cr == [n]
    assert rel_rms(yg, yr) <= TOL
    # the +50 kHz tone is in the stop band: output power must be dominated by the 1 kHz tone
    assert np.mean(np.abs(yg[4096:]) ** 2) < 0.3


def test_filter_spectrum_matches_reference_design():
    import pebblesdr_b200 as pb
    from oracle import ref
    g = pb.CFastFIR()
    g.SetupParameters(300.0, 3000.0, 0.0, 78125.0)
    H = g.GetFilter()
    # reference filter spectrum == its impulse response through ProcessData: feed a unit impulse
    r = ref.CFastFIR()
    r.SetupParameters(300.0, 3000.0, 0.0, 78125.0)
    x = np.zeros(2048, complex)
    x[0] = 1.0
    out = np.zeros(4096, complex)
    k = r.ProcessData(2048, x, out)
    h = np.fft.ifft(H) * 2048  # taps/2048 -> FFT -> H ; ifft gives taps/2048
    assert k == 2048
    assert rel_rms(h[:1025] , out[:1025]) <= 1e-9


@pytest.mark.parametrize("sizes", [[2048] * 16, [1000, 24, 1024, 3000, 1, 1, 2046, 5000, 777], [1023, 1, 1024, 4096],
                                   # 6+ reference blocks per call: the triple-block (4096-point) kernel with 0, 1 and 2
                                   # left-over reference blocks, mixed with calls that only take the 2048-point kernel
                                   [6144, 7168, 8192, 5120, 9216, 10240], [7000, 6500, 8100, 1, 13000, 40000]])
def test_split_stream_counts_and_samples(sizes):
    # fastfir.cpp:299-313 block boundaries: first block after 1024 inputs; any split gives the same stream
    n = sum(sizes)
    x = iq_noise_tones(n, 99, 2e6, [(2e3, 0.3)])
    g, r = _pair()
    yg, cg = _feed(g, x, sizes)
    yr, cr = _feed(r, x, sizes)
    assert cg == cr
    assert len(yg) == len(yr)
    assert rel_rms(yg, yr) <= TOL


def test_empty_and_param_error():
    import pebblesdr_b200 as pb
    g = pb.CFastFIR()
    assert g.SetupParameters(-3000.0, 3000.0, 0.0, 2e6) == 0
    assert g.SetupParameters(-3000.0, 3000.0, 0.0, 2e6) == 1          # unchanged
    assert g.SetupParameters(3000.0, -3000.0, 0.0, 2e6) == 2          # "Filter Parameter error": old filter kept
    out = np.zeros(16, complex)
    assert g.ProcessData(0, np.zeros(0, complex), out) == 0
    x = iq_noise_tones(4096, 5, 2e6, [(1e3, 0.5)])
    from oracle import ref
    r = ref.CFastFIR()
    r.SetupParameters(-3000.0, 3000.0, 0.0, 2e6)
    yg, _ = _feed(g, x, [4096])
    yr, _ = _feed(r, x, [4096])
    assert rel_rms(yg, yr) <= TOL


def test_in_place_device_call_is_refused_host_call_is_not():
    import torch
    import pebblesdr_b200 as pb
    from pebblesdr_b200._lib import lib
    x = iq_noise_tones(4096, 12, 2e6, [(3e3, 0.4)], 0.05)
    f = pb.CFastFIR()
    f.SetupParameters(-5000.0, 5000.0, 0.0, 2e6)
    d = torch.from_numpy(x).cuda()
    assert lib.pebble_fastfir_process_device(f._h, 4096, d.data_ptr(), d.data_ptr(), None) < 0
    assert b"overlap" in lib.pebble_last_error()
    assert lib.pebble_fastfir_process_device(f._h, 4096, d.data_ptr(), d.data_ptr() + 16 * 1000, None) < 0
    g, r = pb.CFastFIR(), pb.CFastFIR()
    for o in (g, r):
        o.SetupParameters(-5000.0, 5000.0, 0.0, 2e6)
    buf, out = x.copy(), np.zeros(4096, complex)
    assert g.ProcessData(4096, buf, buf) == r.ProcessData(4096, x, out) == 4096      # in place through the host call
    assert np.array_equal(buf, out)


def test_state_roundtrip():
    g, _ = _pair()
    x = iq_noise_tones(5000, 3, 2e6, [(1e3, 0.5)])
    out = np.zeros(8192, complex)
    g.ProcessData(2500, x[:2500].copy(), out)
    st = g.GetState()
    k1 = g.ProcessData(2500, x[2500:].copy(), out)
    a = out[:k1].copy()
    g.SetState(st)
    k2 = g.ProcessData(2500, x[2500:].copy(), out)
    assert k1 == k2 and np.array_equal(a, out[:k2])


def test_device_resident_large_linearity():
    # full-size property check: linearity of the filter on device-resident buffers (2^22 samples)
    import torch
    import pebblesdr_b200 as pb
    n = 1 << 22
    gen = torch.Generator(device="cuda").manual_seed(1)
    a = torch.randn(n, 2, device="cuda", dtype=torch.float64, generator=gen)
    b = torch.randn(n, 2, device="cuda", dtype=torch.float64, generator=gen)
    xa, xb = torch.view_as_complex(a), torch.view_as_complex(b)
    outs = []
    for x in (xa, xb, xa + 2.0 * xb):
        f = pb.CFastFIR()
        f.SetupParameters(-3000.0, 3000.0, 0.0, 2e6)
        y = torch.zeros(n, dtype=torch.complex128, device="cuda")
        assert f.ProcessData(n, x.contiguous(), y) == n
        outs.append(y)
    torch.cuda.synchronize()
    err = (outs[2] - (outs[0] + 2.0 * outs[1])).abs().max().item()
    ref = outs[2].abs().max().item()
    assert err <= 1e-12 * max(ref, 1.0)


def test_fused_gain_rides_on_filter_spectrum():
    # the receiver folds the decimation gain (receiver.cpp:935-938) and the fixed AGC gain into the FastFIR call; the
    # fused kernels carry the product on H and must pick up a change between two calls of the same stream
    g, r = _pair()
    n = 1 << 15
    x = iq_noise_tones(2 * n, 5, 2e6, [(1.5e3, 0.4)])
    yr, _ = _feed(r, x, [n, n])
    g.SetFusedScale(2.5, 0.5)
    y0, _ = _feed(g, x[:n], [n])
    g.SetFusedScale(4.0, 3.0)
    y1, _ = _feed(g, x[n:], [n])
    assert rel_rms(y0, 1.25 * yr[:n]) <= TOL
    assert rel_rms(y1, 12.0 * yr[n:]) <= TOL


# ---- long filters (BASELINE config 5: 131072-point overlap-save, 65537 taps; four-step kernels, fastfir_long.cu) ----------
def _long_oracle(N):
    """The reference compiled at 131072/65537 when present, else the C restatement (pinned to it, tests/test_oracle.py)."""
    from oracle import ref, port
    if N == 131072 and ref.long_available():
        return ref.CFastFIRLong()
    return port.CFastFIR(N, N // 2 + 1)


@pytest.mark.parametrize("N,calls", [(32768, [16384 * 5]), (32768, [7, 16377, 40000, 9152]),
                                     (131072, [65536 * 3]), (131072, [65536, 5000, 2 * 65536 - 5000, 4096, 61440]),
                                     # >= 7 blocks in a call: super-blocks (4N-point transforms, 7 blocks each) + the remainder
                                     (32768, [16384 * 17]), (32768, [16384 * 7 + 100, 16384 * 9 - 95, 16384 * 14]),
                                     (131072, [65536 * 9]), (131072, [65536 * 7 + 123, 65536 * 8])])
def test_long_filter_vs_reference(N, calls, monkeypatch):
    import pebblesdr_b200 as pb
    L, n = N // 2, sum(calls)
    if n >= 14 * L:
        monkeypatch.setenv("PEBBLE_LF_PIPE_MIN", "1")              # the longest cases also run the persistent pipeline kernel
    x = iq_noise_tones(n, 5000, 100e6, [(7e6, 0.4), (31e6, 0.5)], 0.1)
    g, r = pb.CFastFIR(N, L + 1), _long_oracle(N)
    assert g.SetupParameters(-20e6, 20e6, 0.0, 100e6) == 0
    r.SetupParameters(-20e6, 20e6, 0.0, 100e6)
    yg, yr, pos = [], [], 0
    for c in calls:
        xi = np.ascontiguousarray(x[pos:pos + c])
        og, orr = np.zeros(c + N, complex), np.zeros(c + N, complex)
        kg, kr = g.ProcessData(c, xi, og), r.ProcessData(c, xi.copy(), orr)
        assert kg == kr                                            # counts and block boundaries exact
        yg.append(og[:kg].copy()); yr.append(orr[:kr].copy())
        pos += c
    yg, yr = np.concatenate(yg), np.concatenate(yr)
    assert len(yg) == (n // L) * L
    assert rel_rms(yg, yr) <= TOL


def test_long_filter_spectrum_and_parameter_error():
    import pebblesdr_b200 as pb
    from oracle import port
    N = 131072
    g = pb.CFastFIR(N, N // 2 + 1)
    assert g.SetupParameters(-20e6, 20e6, 0.0, 100e6) == 0
    assert g.SetupParameters(-20e6, 20e6, 0.0, 100e6) == 1        # unchanged (fastfir.cpp:195-199)
    H = g.GetFilter()
    assert g.SetupParameters(10e6, 60e6, 0.0, 100e6) == 2          # "Filter Parameter error": old filter kept
    assert np.array_equal(g.GetFilter(), H)
    # H = FFT of the reference's taps: the oracle's impulse response through ProcessData is taps * N
    o = port.CFastFIR(N, N // 2 + 1)
    o.SetupParameters(-20e6, 20e6, 0.0, 100e6)
    x = np.zeros(N, complex); x[0] = 1.0
    out = np.zeros(2 * N, complex)
    assert o.ProcessData(N, x, out) == N
    h = np.fft.ifft(H) * N
    assert rel_rms(h[:N // 2 + 1], out[:N // 2 + 1]) <= 1e-9
    pb_mag = np.abs(H)
    assert abs(pb_mag[0] * N - 1.0) < 1e-3 and pb_mag[N // 2] * N < 1e-6   # pass band 1/N (fastfir.cpp:243), dead stop band


def test_long_filter_streams_batch_equals_single_streams_and_split():
    """config 5 shape at reduced length: 6 independent streams sharing one filter, device-resident, two calls."""
    import torch
    import pebblesdr_b200 as pb
    from pebblesdr_b200._lib import lib, check
    N, L, S, n = 131072, 65536, 6, 65536 * 4
    xs = np.stack([iq_noise_tones(n, 5000 + s, 100e6, [((s - 3) * 5e6, 0.4)], 0.1) for s in range(S)])
    singles = []
    for s in range(S):
        f = pb.CFastFIR(N, L + 1)
        f.SetupParameters(-20e6, 20e6, 0.0, 100e6)
        o = np.zeros(n + N, complex)
        k = f.ProcessData(n, np.ascontiguousarray(xs[s]), o)
        assert k == n
        singles.append(o[:k].copy())
    b = pb.CFastFIR(N, L + 1)
    b.SetupParameters(-20e6, 20e6, 0.0, 100e6)
    check(lib.pebble_fastfir_set_streams(b._h, S))
    dx = torch.from_numpy(xs).cuda()
    dy = torch.zeros((S, n), dtype=torch.complex128, device="cuda")
    cut = L + 12345                                                # ragged split: second call starts mid-block
    d1 = dx[:, :cut].contiguous()
    d2 = dx[:, cut:].contiguous()
    k1 = check(lib.pebble_fastfir_process_batch_device(b._h, cut, d1.data_ptr(), cut, dy.data_ptr(), n, None))
    assert k1 == L
    k2 = check(lib.pebble_fastfir_process_batch_device(b._h, n - cut, d2.data_ptr(), n - cut,
                                                       dy.data_ptr() + 16 * k1, n, None))
    assert k1 + k2 == n
    torch.cuda.synchronize()
    got = dy.cpu().numpy()
    for s in range(S):
        assert rel_rms(got[s], singles[s]) <= 1e-13, s              # same kernels, same block grid: ~bit-equal
    # linearity at full length (size-independent property): F(a x1 + b x2) = a F(x1) + b F(x2)
    mix = 0.5 * singles[0] - 2.0 * singles[1]
    f = pb.CFastFIR(N, L + 1)
    f.SetupParameters(-20e6, 20e6, 0.0, 100e6)
    o = np.zeros(n + N, complex)
    f.ProcessData(n, np.ascontiguousarray(0.5 * xs[0] - 2.0 * xs[1]), o)
    assert rel_rms(o[:n], mix) <= 1e-12


@pytest.mark.parametrize("N", [32768, 131072])
def test_long_filter_super_blocks_equal_block_by_block(N):
    """A call of 16 blocks (two 4N-point super-blocks of 7 + two N-point blocks) against the same stream fed one block per
    call (N-point plan only), three batched streams, device-resident; and against a handle created with the super-block plan
    disabled. Same linear convolution: FP64 rounding apart (1e-13)."""
    import os
    import torch
    import pebblesdr_b200 as pb
    from pebblesdr_b200._lib import lib, check
    L, S, nb = N // 2, 3, 16
    n = L * nb
    xs = np.stack([iq_noise_tones(n, 5200 + s, 100e6, [((s - 1) * 9e6, 0.4)], 0.1) for s in range(S)])
    dx = torch.from_numpy(xs).cuda()

    def run(step_blocks, **env):
        os.environ.update(env)                                  # read when the handle is created
        try:
            f = pb.CFastFIR(N, L + 1)
        finally:
            for k in env:
                os.environ.pop(k, None)
        f.SetupParameters(-20e6, 20e6, 0.0, 100e6)
        check(lib.pebble_fastfir_set_streams(f._h, S))
        dy = torch.zeros((S, n), dtype=torch.complex128, device="cuda")
        done = 0
        while done < n:
            c = min(step_blocks * L, n - done)
            d = dx[:, done:done + c].contiguous()
            k = check(lib.pebble_fastfir_process_batch_device(f._h, c, d.data_ptr(), c, dy.data_ptr() + 16 * done, n, None))
            assert k == c
            done += c
        torch.cuda.synchronize()
        return dy.cpu().numpy()

    whole, single, off = run(nb), run(1), run(nb, PEBBLE_LF_NOSUPER="1")
    # the persistent pipeline kernel (lf_pipe_kernel) forced on at this small size, with the smallest and the default ring,
    # and forced off: the same tile functions in a different launch structure (inlined twice, so FMA contraction may differ)
    piped = run(nb, PEBBLE_LF_PIPE_MIN="1")
    piped_small_ring = run(nb, PEBBLE_LF_PIPE_MIN="1", PEBBLE_LF_RING="4" if N == 131072 else "6")
    triples = whole                                            # the default: launch triples
    piped_single = run(1, PEBBLE_LF_PIPE_MIN="1")
    for s in range(S):
        assert rel_rms(whole[s], single[s]) <= 1e-13, s
        assert rel_rms(whole[s], off[s]) <= 1e-13, s
        assert rel_rms(single[s], off[s]) <= 1e-13, s
        assert rel_rms(piped[s], triples[s]) <= 1e-14, s
        assert rel_rms(piped_small_ring[s], piped[s]) <= 1e-14, s
        assert rel_rms(piped_single[s], single[s]) <= 1e-13, s


def test_long_filter_small_calls_and_state_round_trip():
    """n = 0, calls shorter than a block, and a stream continued by a fresh object restarted from the saved state."""
    import pebblesdr_b200 as pb
    N, L = 32768, 16384
    x = iq_noise_tones(3 * L + 700, 5100, 100e6, [(3e6, 0.4)], 0.1)
    a = pb.CFastFIR(N, L + 1)
    a.SetupParameters(-20e6, 20e6, 0.0, 100e6)
    whole = np.zeros(4 * L, complex)
    kw = a.ProcessData(len(x), x, whole)
    assert kw == 3 * L
    b = pb.CFastFIR(N, L + 1)
    b.SetupParameters(-20e6, 20e6, 0.0, 100e6)
    out = np.zeros(2 * L, complex)
    assert b.ProcessData(0, x[:0], out) == 0                               # fastfir.cpp:287-288
    assert b.ProcessData(100, np.ascontiguousarray(x[:100]), out) == 0     # fill only
    assert b.OutCount(L - 100) == L and b.OutCount(L - 101) == 0
    got = []
    k = b.ProcessData(L + 5, np.ascontiguousarray(x[100:L + 105]), out)
    assert k == L
    got.append(out[:k].copy())
    st = b.GetState()
    c = pb.CFastFIR(N, L + 1)                                              # fresh object, same filter, saved state
    c.SetupParameters(-20e6, 20e6, 0.0, 100e6)
    c.SetState(st)
    rest = np.ascontiguousarray(x[L + 105:])
    out2 = np.zeros(len(rest) + N, complex)
    k2 = c.ProcessData(len(rest), rest, out2)
    assert k + k2 == kw
    got.append(out2[:k2].copy())
    assert rel_rms(np.concatenate(got), whole[:kw]) <= 1e-13
