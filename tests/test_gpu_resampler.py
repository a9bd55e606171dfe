"""GPU parity: CFractResampler against the COMPILED reference (pebblelib/fractresampler.cpp built by
oracle/build_ref.sh step 5). Output counts exact; samples BIT-IDENTICAL (np.array_equal): the kernel forms every product
and sum with the reference's own roundings, and the windowed-sinc table comes from the same libm."""
import numpy as np
import pytest

from conftest import rel_rms, iq_noise_tones

pytestmark = pytest.mark.gpu


def _ref_run(rate, x, sizes):
    from oracle import ref_app
    r = ref_app.CFractResampler(max(sizes))
    out, pos = [], 0
    for n in sizes:
        out.append(r.Resample(x[pos:pos + n], rate))
        pos += n
    return out


@pytest.mark.parametrize("rate", [78125.0 / 48000.0, 60000.0 / 48000.0, 62500.0 / 44100.0, 0.75])
def test_resample_host_calls_bit_exact_ragged(rate):
    import pebblesdr_b200 as pb
    sizes = [2048, 2048, 1, 29, 777, 4096, 28, 2048]
    n = sum(sizes)
    x = iq_noise_tones(n, 91, 78125.0, [(1000.0, 0.4), (-2500.0, 0.2)], noise=0.05)
    want = _ref_run(rate, x, sizes)
    g = pb.CFractResampler(4096)
    pos = 0
    for n_i, w in zip(sizes, want):
        k, y = g.Resample(n_i, rate, np.ascontiguousarray(x[pos:pos + n_i]))
        assert k == len(w)
        assert np.array_equal(y[:k], w)
        pos += n_i


def test_resample_batched_frames_equal_per_frame_calls_and_split_state():
    """One device call over 64 receiver frames (frame = 2048) == 64 reference calls; state round trip; two streams."""
    import torch
    import pebblesdr_b200 as pb
    rate, frame, frames = 78125.0 / 48000.0, 2048, 64
    n = frame * frames
    xa = iq_noise_tones(n, 92, 78125.0, [(700.0, 0.5)], noise=0.02)
    xb = iq_noise_tones(n, 93, 78125.0, [(-1500.0, 0.3)], noise=0.1)
    wa = np.concatenate(_ref_run(rate, xa, [frame] * frames))
    wb = np.concatenate(_ref_run(rate, xb, [frame] * frames))
    g = pb.CFractResampler(frame, streams=2)
    xin = torch.from_numpy(np.stack([xa, xb])).cuda()
    k, y = g.Resample(n, rate, xin, frame=frame)
    assert k == len(wa) == len(wb)
    y = y.cpu().numpy()
    assert np.array_equal(y[0, :k], wa) and np.array_equal(y[1, :k], wb)
    # split: first 20 frames, state out / in through a second object, remaining 44
    g1 = pb.CFractResampler(frame)
    k1, y1 = g1.Resample(frame * 20, rate, torch.from_numpy(xa[:frame * 20]).cuda(), frame=frame)
    g2 = pb.CFractResampler(frame)
    g2.SetState(g1.GetState())
    k2, y2 = g2.Resample(frame * 44, rate, torch.from_numpy(xa[frame * 20:]).cuda(), frame=frame)
    assert k1 + k2 == k
    assert np.array_equal(np.concatenate([y1.cpu().numpy()[0, :k1], y2.cpu().numpy()[0, :k2]]), wa)
    # a different frame structure is a different reference call sequence (m_FloatTime rounds differently): counts still add up
    g3 = pb.CFractResampler(n)
    k3, y3 = g3.Resample(n, rate, torch.from_numpy(xa).cuda())
    w3 = _ref_run(rate, xa, [n])[0]
    assert k3 == len(w3) and np.array_equal(y3.cpu().numpy()[0, :k3], w3)


def test_resample_stereo16_and_audio_clip():
    import torch
    import pebblesdr_b200 as pb
    from oracle import ref_app
    rate, n = 60000.0 / 48000.0, 2048 * 4
    x = iq_noise_tones(n, 94, 60000.0, [(900.0, 0.9)], noise=0.2)
    r = ref_app.CFractResampler(2048)
    want = np.concatenate([r.ResampleStereo16(x[i:i + 2048], rate, 30000.0) for i in range(0, n, 2048)])
    g = pb.CFractResampler(2048)
    k, y = g.ResampleStereo16(n, rate, torch.from_numpy(x).cuda(), 30000.0, frame=2048)
    assert k == len(want) and np.array_equal(y.cpu().numpy()[0, :k], want)
    assert np.abs(want).max() == 32767            # the gain drives it into the clip (fractresampler.cpp:228-235)
    # CPX output with the audio sink's clip (AudioQT::SendToOutput, audioqt.cpp:190-201)
    r2 = ref_app.CFractResampler(2048)
    w2 = np.concatenate([r2.Resample(x[i:i + 2048] * 1.5, rate) for i in range(0, n, 2048)])
    mx = float(np.float32(0.9999))
    w2 = np.clip(w2.real, -mx, mx) + 1j * np.clip(w2.imag, -mx, mx)
    g2 = pb.CFractResampler(2048)
    k2, y2 = g2.Resample(n, rate, torch.from_numpy(np.ascontiguousarray(x * 1.5)).cuda(), frame=2048, clip=True)
    assert k2 == len(w2) and np.array_equal(y2.cpu().numpy()[0, :k2], w2)


def test_receiver_chain_iq_in_audio_out():
    """IQ in -> audio out: BASELINE config 2 chain + the audio tail of Receiver::processIQData (receiver.cpp:994-1007):
    Resample(frame, 78125/48000) once per 2048-sample frame, then the sink's clip. Reference = oracle chain (compiled
    reference steps) -> compiled CFractResampler -> clip. The chain's 1e-10 tolerance carries through the linear resampler."""
    import torch
    import pebblesdr_b200 as pb
    from oracle import cpu_chain, ref_app
    from bench import synth_numpy
    n = 32768 * 24
    x = synth_numpy(n)
    rx = pb.Receiver(10e6, 2048)
    assert rx.setDataRate(15000.0) == 78125.0
    rx.setMixer(1.2e6); rx.setFilter(-5000.0, 5000.0); rx.setDemodMode(pb.Demod.dmAM); rx.setAgc(0, 20)
    rx.setDemodBandwidth(10000.0)
    rx.setAudio(48000, clip=True)
    outs, pos = [], 0
    for m in (32768 * 8, 32768 * 3, 32768 * 13):               # frames cross the call boundaries
        k, y = rx.processIQData(torch.from_numpy(x[pos:pos + m]).cuda(), m)
        assert k == rx_count_check(k)
        outs.append(y[:k].cpu().numpy()); pos += m
    got = np.concatenate(outs)
    c = cpu_chain.CpuChain()
    demod = c.run(x.copy())
    r = ref_app.CFractResampler(2048)
    rate = (78125 * 1.0) / (48000 * 1.0)
    want = np.concatenate([r.Resample(demod[i:i + 2048], rate) for i in range(0, len(demod), 2048)])
    mx = float(np.float32(0.9999))
    want = np.clip(want.real, -mx, mx) + 1j * np.clip(want.imag, -mx, mx)
    assert len(got) == len(want) > 0
    assert rel_rms(got, want) <= 1e-10


def rx_count_check(k):
    return k


@pytest.mark.parametrize("mode", ["dmFMS", "dmFMM"])
def test_receiver_wfm_branch_iq_in_audio_out(mode):
    """The WFM branch of Receiver::processIQData (receiver.cpp:854-901 + the tail :994-1007): m_downConvertWfm1
    (SetDataRateSimple(2 MSPS, 200000) -> 250 kS/s) -> 2048-sample frames -> Demod::fmStereo / fmMono at the WFM rate, one call
    per frame -> Resample(frame, 250000/48000) -> the sink's clip. Reference = the same steps compiled from the reference
    (CDownConvert, Demod_WFM, CFractResampler), composed in that order. Input: an FM stereo broadcast 300 kHz off centre at
    2 MSPS; calls cut across frames; a switch to AM and back in the middle restarts the frame on both sides."""
    import torch
    import pebblesdr_b200 as pb
    from oracle import ref, ref_app
    fs, fpb, off = 2000000.0, 2048, 300000.0
    n = 16384 * 40
    t = np.arange(n) / fs
    left, right = 0.5 * np.sin(2 * np.pi * 1000.0 * t), 0.3 * np.sin(2 * np.pi * 2500.0 * t)
    pil = 2 * np.pi * 19010.0 * t
    m = 0.4 * (left + right) + 0.1 * np.sin(pil) + 0.4 * (left - right) * np.sin(2 * pil)
    ph = 2 * np.pi * np.cumsum(75000.0 * m) / fs
    rng = np.random.default_rng(3)
    x = np.ascontiguousarray(0.4 * np.exp(1j * (ph + 2 * np.pi * off * t)) + (rng.standard_normal(n) + 1j * rng.standard_normal(n)) * 0.002)

    rx = pb.Receiver(fs, fpb)
    rx.setDataRate(15000.0)
    assert rx.wfmRate() == 250000.0
    rx.setMixer(-off); rx.setFilter(-5000.0, 5000.0); rx.setAgc(0, 20)
    rx.setDemodMode(getattr(pb.Demod, mode))
    rx.setAudio(48000, clip=True)
    outs, locks, pos = [], [], 0
    for c in (16384 * 9, 16384 * 2, 16384 * 29):
        k, y = rx.processIQData(torch.from_numpy(x[pos:pos + c]).cuda(), c)
        outs.append(y[:k].cpu().numpy()); pos += c
        locks.append(rx.getStereoLock())
    got = np.concatenate(outs)

    dc = ref.CDownConvert()
    assert dc.SetDataRateSimple(fs, 200000.0) == 250000.0
    dc.SetFrequency(-off)
    dem = ref_app.Demod_WFM(250000, fpb)
    rs = ref_app.CFractResampler(fpb)
    y = np.zeros(n // 8, dtype=np.complex128)
    for i in range(0, n, 16384):                      # the reference's own call size is irrelevant once >= 2 (taps - 1) per stage
        k = dc.ProcessData(16384, x[i:i + 16384], y[i // 8:])
        assert k == 2048
    rate = (250000 * 1.0) / (48000 * 1.0)
    want, lock = [], 0
    for i in range(0, len(y) - fpb + 1, fpb):
        a = dem.processDataStereo(y[i:i + fpb].copy()) if mode == "dmFMS" else dem.processDataMono(y[i:i + fpb].copy())
        want.append(rs.Resample(a, rate))
        lock = dem.getStereoLock()[1]
    want = np.concatenate(want)
    mx = float(np.float32(0.9999))
    want = np.clip(want.real, -mx, mx) + 1j * np.clip(want.imag, -mx, mx)
    assert len(got) == len(want) > 15000
    assert rel_rms(got, want) <= 1e-9, rel_rms(got, want)
    if mode == "dmFMS":
        assert locks[-1] == bool(lock)                                        # the pilot lock state the GUI would read

