"""bench_channelizer.py — BASELINE config 4 leg of bench.py (`python bench.py --workload channelizer [--gpus N]`):
1024 independent NBFM/AM receivers over one synthetic 61.44 MSPS capture, channels sharded across the ranks
(strong scaling: total work fixed), ONE gather of the audio onto rank 0 over NCCL per step.

A step = every channel processes one batch of 2^LOG2N capture samples (default 2^24 = 256 MiB > the 126 MB L2).
metric value = channel-samples/s = capture samples x channels / time, in complex MS/s (SURVEY 8(d) config 4:
"input rate x channels"). The path is FP64-bound, not HBM-bound (every channel mixes the full-rate stream; the
capture is shared through L2): the roofline object reports the HBM view the contract asks for AND the FP64 view."""
import json
import math
import os
import time

FS, MAX_BW = 61.44e6, 15000.0
ALG_BYTES_PER_INPUT_SAMPLE = 32.0          # 16 B read + 1024 x 16/1024 B written
FLOP_PER_CHANNEL_SAMPLE = 22.0             # per-channel front end: ~22 flop per input sample per channel
# polyphase-DFT front end, per capture sample: mix 6 + oscillator step 6 + 25 taps x 2 components x 2 flop (polyphase FIR) +
# 5 N log2 N / N = 50 (1024-point DFT) + 6/1024 x channels (output rotation) + FastFIR 2 x 5 x 2048 x 11 / 1024 per decimated
# sample x channels / 1024 ~ 220: every channel's band-pass runs at 1/1024 of the capture rate, so it adds 220 per capture sample
FLOP_PER_INPUT_SAMPLE_POLY = 6 + 6 + 100 + 50 + 6 + 220


def synth_capture(n, device, nch_total, seed=4567):
    """Sum of 1024 carriers (even k NBFM 5 kHz dev / 1 kHz tone, odd k AM m=0.5) at -30.69 MHz + k*60 kHz, amp 1/64,
    plus Gaussian noise. Synthesised with a 1024-point inverse DFT per 1024-sample block would be exact only for
    unmodulated carriers, so the sum is built directly in chunks (setup cost, untimed)."""
    import torch
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    x = torch.empty(n, dtype=torch.complex128, device=device)
    xr = torch.view_as_real(x)
    blk = 1 << 16
    k = torch.arange(nch_total, dtype=torch.float64, device=device)
    fc = (-30.69e6 + 60e3 * k) / FS                      # cycles per sample
    even = (torch.arange(nch_total, device=device) % 2 == 0)
    for i in range(0, n, blk):
        m = min(blk, n - i)
        t = torch.arange(i, i + m, dtype=torch.float64, device=device)
        tone = 2 * math.pi * torch.remainder((1000.0 / FS) * t, 1.0)
        acc = torch.randn(m, 2, generator=g, dtype=torch.float64, device=device) * 0.002
        for c0 in range(0, nch_total, 128):              # 128 carriers at a time: [128, m] temporaries
            f = fc[c0:c0 + 128, None]
            ph = 2 * math.pi * torch.remainder(f * t[None, :], 1.0)
            ev = even[c0:c0 + 128, None]
            ph = ph + torch.where(ev, 5.0 * torch.sin(tone)[None, :], torch.zeros((), dtype=torch.float64, device=device))
            amp = torch.where(ev, torch.ones((), dtype=torch.float64, device=device), 1 + 0.5 * torch.cos(tone)[None, :]) / 64.0
            acc[:, 0] += (amp * torch.cos(ph)).sum(0)
            acc[:, 1] += (amp * torch.sin(ph)).sum(0)
        xr[i:i + m] = acc
    return x


def cpu_baseline(args, nch_total):
    """The same receivers on the host cores, one channel per thread, bounded sample. kind "reference": every step is the
    compiled reference (oracle/cpu_chain.py ChannelChain: CDownConvert built with MAX_DECSTAGES 12 because the stock build
    overflows on this 10-stage chain); kind "port": the C restatement when oracle/_ref is absent."""
    import threading
    import numpy as np
    from oracle import port, cpu_chain
    threads = cpu_chain.host_threads()
    m = 65536 * 16
    rng = np.random.default_rng(1)
    src = ((rng.standard_normal(m) + 1j * rng.standard_normal(m)) * 0.05).astype(np.complex128)
    real = cpu_chain.channel_chain_available()
    rx = []
    for i in range(threads):
        if real:
            rx.append(cpu_chain.ChannelChain(FS, MAX_BW, -30.69e6 + 60e3 * i, 2 if i % 2 == 0 else 0))
        else:
            o = port.Receiver(FS, 2048, MAX_BW)
            o.config(-30.69e6 + 60e3 * i, -7500.0, 7500.0, 2 if i % 2 == 0 else 0, 0, 20)
            rx.append(o)
    busy = [0.0] * threads
    start = threading.Barrier(threads)

    def work(i):
        start.wait()
        t0 = time.perf_counter()
        if real:
            rx[i].run(src)
        else:
            for p in range(0, m, 65536):
                rx[i].processIQData(src[p:p + 65536], 65536)
        busy[i] = time.perf_counter() - t0

    ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return {"value": m * threads / max(busy) / 1e6, "unit": "MS/s", "cores": threads, "kind": "reference" if real else "port",
            "sample": "%d receivers (one channel per host thread) x %d capture samples, 61.44 MSPS 10-stage chain + FastFIR + "
                      "NBFM/AM demod; %s" % (threads, m, "compiled reference steps (CDownConvert with MAX_DECSTAGES 12, CFastFIR + "
                      "FFTW 3.3.4, AGC, Demod_NFM / Demod_AM)" if real else "C restatement oracle/pebble_oracle.c")}


PARITY_CHANNELS = [0, 1, 511, 512, 1022, 1023, 300, 777]
PARITY_LOG2 = 22          # capture samples replayed through the reference chain per checked channel
LOCK = 1024               # NBFM lock-time allowance (output samples), see tests/test_gpu_channelizer.py


def channel_parity(x_head, rows, freqs, channels):
    """rows[c]: the first len(x_head) >> 10 demodulated samples channel c produced in the bench's own run (first step of
    the stream) vs the reference chain (oracle/cpu_chain.py ChannelChain, compiled reference) on the same capture samples.
    AM channels <= 1e-10; NBFM (float PLL state) <= 1e-4 after the lock-time allowance."""
    import numpy as np
    from oracle import cpu_chain
    if not cpu_chain.channel_chain_available():
        return {"rel_rms": None, "oracle": "unavailable (oracle/_ref not built)"}
    worst_am = worst_fm = 0.0
    for c in channels:
        mode = 2 if c % 2 == 0 else 0
        want = cpu_chain.ChannelChain(FS, MAX_BW, freqs[c], mode).run(x_head)
        got = rows[c][:len(want)]
        lo = LOCK if mode == 2 else 0
        den = float(np.sqrt(np.mean(np.abs(want[lo:]) ** 2)))
        err = float(np.sqrt(np.mean(np.abs(got[lo:] - want[lo:]) ** 2))) / (den if den > 0 else 1.0)
        if mode == 2:
            worst_fm = max(worst_fm, err)
        else:
            worst_am = max(worst_am, err)
    return {"rel_rms": worst_am, "rel_rms_nbfm_after_lock": worst_fm, "tolerance": 1e-10, "tolerance_nbfm": 1e-4,
            "channels": list(channels), "samples": int(len(x_head) >> 10), "oracle": "reference",
            "what": "worst channel: first %d demodulated samples of the run's first step vs the compiled reference chain "
                    "(CDownConvert MAX_DECSTAGES 12 -> CFastFIR -> AGC_OFF -> Demod_AM / Demod_NFM) on the same capture"
                    % (len(x_head) >> 10)}


def measure(args, bench, torch, dist, dev, world, rank, steps, warmup, cpu_baseline=False, e2e=False):
    """Times the sharded channelizer on an already initialised process group. Returns the result dict on rank 0."""
    import pebblesdr_b200 as pb
    from pebblesdr_b200 import lib
    from pebblesdr_b200.dist import shard_range, gather_rows, max_over_ranks, channel_frequencies

    nch_total = args.channels
    log2n = args.log2n if (args.log2n_set and args.workload == "channelizer") else 24
    n = 1 << log2n
    K, W = steps, max(warmup, 1)
    c0, c1 = shard_range(nch_total, world, rank)
    freqs = channel_frequencies(nch_total)
    x = synth_capture(n, dev, nch_total)                 # replicated input: every rank synthesises the same capture
    cz = pb.Channelizer(c1 - c0, FS, MAX_BW)
    assert cz.outRate == 60000.0 and cz.numStages == 10
    cz.setFilter(-7500.0, 7500.0)
    for c in range(c0, c1):
        cz.setFrequency(c - c0, freqs[c])
        cz.setDemodMode(c - c0, 2 if c % 2 == 0 else 0)  # even NBFM (PLL), odd AM
    nd = cz.outCount(n)
    out = torch.zeros(c1 - c0, nd, dtype=torch.complex128, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        cz.process(x, n, out)
        return gather_rows(out, nch_total, dst=0)        # the one collective of the path (audio -> rank 0)

    first_rows = None
    for w in range(W):
        full = step()
        if w == 0 and rank == 0:                         # the stream's first step: what the parity check replays
            npar = min(nd, (1 << PARITY_LOG2) >> 10)
            chans = [c for c in PARITY_CHANNELS if c < nch_total]
            first_rows = {c: full[c, :npar].cpu().numpy() for c in chans}
    barrier()
    clocks = bench.ClockSampler(torch.cuda.current_device())
    if rank == 0:
        clocks.start()
        time.sleep(0.3)
    l0 = lib.pebble_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(K):
        full = step()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1), dev)
    launches = lib.pebble_launch_count() - l0
    front_path = cz.frontPath()
    # the two halves separately: compute only, and the gather only (both device-timed, max over ranks)
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    g0.record()
    for _ in range(K):
        cz.process(x, n, out)
    g1.record()
    barrier()
    ms_compute = max_over_ranks(g0.elapsed_time(g1), dev)
    g0.record()
    for _ in range(K):
        full = gather_rows(out, nch_total, dst=0)
    g1.record()
    barrier()
    ms_gather = max_over_ranks(g0.elapsed_time(g1), dev) if world > 1 else 0.0
    clk = clocks.stop() if rank == 0 else None
    if rank == 0:
        chk = float(torch.view_as_real(full).abs().sum().item())
        assert full.shape == (nch_total, nd) and chk > 0 and chk == chk

    ms_e2e = None
    if e2e:
        # ---- end to end: capture in pinned host memory on every rank, audio back in rank 0's pinned host memory ----
        pin_in = torch.empty(n, dtype=torch.complex128, pin_memory=True)
        pin_in.copy_(x)
        pin_out = torch.zeros(nch_total, nd, dtype=torch.complex128, pin_memory=True) if rank == 0 else None
        xin = torch.empty_like(x)
        for it in range(2 + K):
            if it == 2:
                f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                barrier()
                f0.record()
            xin.copy_(pin_in, non_blocking=True)
            cz.process(xin, n, out)
            g = gather_rows(out, nch_total, dst=0)
            if rank == 0:
                pin_out.copy_(g, non_blocking=True)
        f1.record()
        barrier()
        ms_e2e = max_over_ranks(f0.elapsed_time(f1), dev)

    if rank != 0:
        return None
    peak, peak_src = bench.measured_peak()
    fpk, fpk_src = bench.fp64_peak()
    cs = float(n) * nch_total * K / (ms * 1e-3)             # channel-samples per second
    in_rate = float(n) * K / (ms * 1e-3)
    parity = channel_parity(x[:1 << PARITY_LOG2].cpu().numpy(), first_rows, freqs, sorted(first_rows)) if first_rows else None
    flop = FLOP_PER_INPUT_SAMPLE_POLY if front_path == 1 else FLOP_PER_CHANNEL_SAMPLE * nch_total / world
    res = {
        "workload": "BASELINE config 4: batched channelizer, %d channels (even NBFM PLL, odd AM) on the 60 kHz grid of a synthetic "
                    "61.44 MSPS capture, CDownConvert chain (/1024) -> CFastFIR +-7.5 kHz -> AGC_OFF -> demod, channels sharded "
                    "over %d GPU(s), one gather of the audio rows per step" % (nch_total, world),
        "metric": "complex MS/s through the batched channelizer (capture rate x channels)", "value": cs / 1e6, "unit": "MS/s",
        "channels": nch_total, "channels_per_gpu": c1 - c0, "n_gpus": world, "steps": K, "warmup": W, "samples_per_step": n,
        "ms_per_step": ms / K, "compute_ms": ms_compute / K, "gather_ms": ms_gather / K,
        "capture_MSps": in_rate / 1e6, "realtime_factor": in_rate / FS,
        "scaling": "strong", "collective": "all_gather_into_tensor (NCCL)" if world > 1 else "none (1 GPU)",
        "front_end": "polyphase-DFT (one mix + 1024-branch polyphase FIR + 1024-point DFT for all channels)" if front_path == 1
                     else "one down-converter per channel",
        "l2": "capture (%d MiB/step) larger than the 126 MB L2; no flush" % ((16 << log2n) >> 20),
        "parity": parity,
        "roofline": {"hbm": {"achieved": ALG_BYTES_PER_INPUT_SAMPLE * in_rate / 1e9, "peak": peak, "unit": "GB/s",
                             "frac": ALG_BYTES_PER_INPUT_SAMPLE * in_rate / 1e9 / peak, "peak_source": peak_src,
                             "algorithmic_bytes_per_step": ALG_BYTES_PER_INPUT_SAMPLE * n, "per": "GPU (every rank reads the capture)"},
                     "fp64": {"achieved": flop * in_rate / 1e12, "peak": fpk, "unit": "TFLOP/s", "frac": flop * in_rate / 1e12 / fpk,
                              "flop_per_capture_sample": flop, "peak_source": fpk_src},
                     "note": "whole step; the front end is ~%d flop per capture sample, the step is bound by the serial per-channel "
                             "demodulator recurrences (PLL / DC block), see DESIGN.md section 4.3" % int(flop)},
        "gpu_launches": int(launches), "clocks": clk,
    }
    if ms_e2e is not None:
        res["e2e"] = {"value": float(n) * nch_total * K / (ms_e2e * 1e-3) / 1e6, "unit": "MS/s",
                      "h2d_bytes_per_step": 16 * n * world, "d2h_bytes_per_step": 16 * nd * nch_total, "ms_per_step": ms_e2e / K}
    return res


def main(args, bench):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    r = measure(args, bench, torch, dist, dev, world, rank, args.steps, args.warmup, e2e=True)
    if rank == 0:
        line = {
            "metric": r["metric"], "value": r["value"], "unit": r["unit"], "n_gpus": world, "steps": r["steps"], "warmup": r["warmup"],
            "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": r["workload"], "samples_per_step": r["samples_per_step"], "channels": r["channels"],
                       "capture_MSps": r["capture_MSps"], "realtime_factor": r["realtime_factor"], "l2": r["l2"],
                       "front_end": r["front_end"], "parallelism": "channels sharded %d-way, replicated input" % world},
            "e2e": r.get("e2e"),
            "roofline": dict(r["roofline"]["hbm"], bound="hbm", traffic=None, fp64=r["roofline"]["fp64"], note=r["roofline"]["note"],
                             kernel="poly_fir_kernel + poly_dft_kernel + fastfir (batched) + chan_demod_kernel + chan_cfir_kernel"),
            "parity": r["parity"], "compute_ms": r["compute_ms"], "gather_ms": r["gather_ms"],
            "gpu_launches": r["gpu_launches"], "clocks": r["clocks"],
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, args.channels)
        bench.emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0
