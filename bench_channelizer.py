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
FLOP_PER_CHANNEL_SAMPLE = 22.0             # ~22 flop per input sample per channel (osc + mix + CIC3 ...)
FP64_NOMINAL_TFLOPS = 37.0                 # no measured FP64 peak in MEASURED_PEAKS.json; nominal B200, stated as such


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
    """The same receivers on the host cores with the C restatement (the reference cannot run the 10-stage chain:
    MAX_DECSTAGES overflow, finding F8): one channel per thread, bounded sample."""
    import threading
    import numpy as np
    from oracle import port, cpu_chain
    threads = cpu_chain.host_threads()
    m = 65536 * 16
    rng = np.random.default_rng(1)
    src = ((rng.standard_normal(m) + 1j * rng.standard_normal(m)) * 0.05).astype(np.complex128)
    rx = []
    for i in range(threads):
        o = port.Receiver(FS, 2048, MAX_BW)
        o.config(-30.69e6 + 60e3 * i, -7500.0, 7500.0, 2 if i % 2 == 0 else 0, 0, 20)
        rx.append(o)
    busy = [0.0] * threads
    start = threading.Barrier(threads)

    def work(i):
        start.wait()
        t0 = time.perf_counter()
        for p in range(0, m, 65536):
            rx[i].processIQData(src[p:p + 65536], 65536)
        busy[i] = time.perf_counter() - t0

    ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return {"value": m * threads / max(busy) / 1e6, "unit": "MS/s", "cores": threads, "kind": "port",
            "sample": "%d receivers (one channel per host thread) x %d capture samples, 61.44 MSPS 10-stage chain + "
                      "FastFIR + NBFM/AM demod, C restatement oracle/pebble_oracle.c (the reference overflows "
                      "MAX_DECSTAGES on this chain, finding F8)" % (threads, m)}


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
    import pebblesdr_b200 as pb
    from pebblesdr_b200 import lib
    from pebblesdr_b200.dist import shard_range, gather_rows, max_over_ranks, channel_frequencies

    nch_total = args.channels
    log2n = args.log2n if args.log2n_set else 24
    n = 1 << log2n
    K, W = args.steps, args.warmup
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

    for _ in range(W):
        full = step()
    barrier()
    clocks = bench.ClockSampler(local)
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
    clk = clocks.stop() if rank == 0 else None
    if rank == 0:
        chk = float(torch.view_as_real(full).abs().sum().item())
        assert full.shape == (nch_total, nd) and chk > 0 and chk == chk

    # ---- end to end: capture in pinned host memory on every rank, audio back in rank 0's pinned host memory ----
    pin_in = torch.empty(n, dtype=torch.complex128, pin_memory=True)
    pin_in.copy_(x)
    pin_out = torch.zeros(nch_total, nd, dtype=torch.complex128, pin_memory=True) if rank == 0 else None
    xin = torch.empty_like(x)
    for _ in range(2):
        xin.copy_(pin_in, non_blocking=True)
        cz.process(xin, n, out)
        g = gather_rows(out, nch_total, dst=0)
        if rank == 0:
            pin_out.copy_(g, non_blocking=True)
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    f0.record()
    for _ in range(K):
        xin.copy_(pin_in, non_blocking=True)
        cz.process(xin, n, out)
        g = gather_rows(out, nch_total, dst=0)
        if rank == 0:
            pin_out.copy_(g, non_blocking=True)
    f1.record()
    barrier()
    ms_e2e = max_over_ranks(f0.elapsed_time(f1), dev)

    if rank == 0:
        peak, peak_src = bench.measured_peak()
        cs = float(n) * nch_total * K / (ms * 1e-3)             # channel-samples per second
        in_rate = float(n) * K / (ms * 1e-3)
        line = {
            "metric": "complex MS/s through the batched channelizer (capture rate x channels)", "value": cs / 1e6,
            "unit": "MS/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "BASELINE config 4: batched channelizer, %d channels (even NBFM PLL, odd AM) from a "
                                   "synthetic 61.44 MSPS capture, 10-stage CDownConvert -> CFastFIR +-7.5 kHz -> demod, "
                                   "channels sharded over %d GPU(s), one NCCL gather of the audio per step" % (nch_total, world),
                       "samples_per_step": n, "channels": nch_total, "capture_MSps": in_rate / 1e6,
                       "realtime_factor": in_rate / FS,
                       "l2": "capture (%d MiB/step) larger than the 126 MB L2; no flush" % ((16 << log2n) >> 20),
                       "parallelism": "channels sharded %d-way, replicated input" % world},
            "e2e": {"value": float(n) * nch_total * K / (ms_e2e * 1e-3) / 1e6, "unit": "MS/s",
                    "h2d_bytes_per_step": 16 * n * world, "d2h_bytes_per_step": 16 * nd * nch_total,
                    "ms_per_step": ms_e2e / K},
            "roofline": {"bound": "hbm", "kernel": "dc_cic_kernel batched (blockIdx = segment x channel)",
                         "achieved": ALG_BYTES_PER_INPUT_SAMPLE * in_rate / world / 1e9 * world, "peak": peak * world,
                         "unit": "GB/s", "frac": ALG_BYTES_PER_INPUT_SAMPLE * in_rate / 1e9 / (peak * world),
                         "traffic": None, "peak_source": peak_src,
                         "note": "FP64-arithmetic-bound by construction (SURVEY 8d config 4); HBM fraction is necessarily small",
                         "fp64": {"achieved_tflops": FLOP_PER_CHANNEL_SAMPLE * cs / 1e12,
                                  "nominal_peak_tflops": FP64_NOMINAL_TFLOPS * world,
                                  "frac_of_nominal": FLOP_PER_CHANNEL_SAMPLE * cs / 1e12 / (FP64_NOMINAL_TFLOPS * world)}},
            "gpu_launches": int(launches), "clocks": clk,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, nch_total)
        bench.emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0
