"""bench_longfir.py — `python bench.py --workload longfir [--gpus N]`: BASELINE config 5, the long-filter stress:
CFastFIR at CONV_FFT_SIZE/CONV_FIR_SIZE 131072/65537 ("65536-tap"), SetupParameters(-20e6, 20e6, 0, 100e6), 64 concurrent
independent streams x 2^22 samples of 100 MSPS IQ. Streams shard across ranks (64 / world each, no
data-path collective: SURVEY 8e), so `scaling` is "strong". Same JSON contract as the chain line.

Algorithmic bytes: 32 B per input sample (16 read + 16 written). The four-step transform's scratch round trips are kept
in L2 by chunking (fastfir_long.cu); whatever still reaches DRAM counts against `achieved`, not the algorithmic figure.
FP64: 2 x 5 N log2 N = 170 N flop per L = N/2 outputs -> 340 flop per sample block by block; a call of >= 7 blocks runs
super-blocks (one 4N-point transform per 7 blocks, fastfir_long.cu): 2 x 5 x 4N x 19 / 7L = 217 flop per sample, and the
figure reported is the one of the transforms actually run (radix-8/16 butterflies need fewer)."""
import os
import time

LF_N, LF_FIR, LF_STREAMS = 131072, 65537, 64
ALG_BYTES = 32.0


def flop_per_sample(n):
    """5 N log2 N convention over the transforms a call of n samples per stream runs (forward + inverse)."""
    L = LF_N // 2
    blocks = n // L
    sup, rem = divmod(blocks, 7)
    return (sup * 10.0 * (4 * LF_N) * 19 + rem * 10.0 * LF_N * 17) / max(blocks * L, 1)


def _cpu_longfir(args):
    """The reference CFastFIR compiled at 131072/65537 (oracle/_ref/libpebbleref_long.so), one stream per host thread."""
    import threading
    import numpy as np
    from oracle import ref, port, cpu_chain
    threads = cpu_chain.host_threads()
    real = ref.long_available()
    m = 1 << 20
    rng = np.random.default_rng(5)
    src = ((rng.standard_normal(m) + 1j * rng.standard_normal(m)) * 0.1).astype(np.complex128)
    objs = []
    for _ in range(threads):
        o = ref.CFastFIRLong() if real else port.CFastFIR(LF_N, LF_FIR)
        o.SetupParameters(-20e6, 20e6, 0.0, 100e6)
        objs.append(o)
    passes = max(1, int(args.cpu_seconds * (12e6 if real else 2e6) / m))
    busy = [0.0] * threads
    gate = threading.Barrier(threads)

    def work(i):
        out = np.zeros(m + LF_N, dtype=np.complex128)
        gate.wait()
        t0 = time.perf_counter()
        for _p in range(passes):
            for q in range(0, m, 1 << 18):
                objs[i].ProcessData(1 << 18, src[q:q + (1 << 18)], out)
        busy[i] = time.perf_counter() - t0

    ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return {"value": m * passes * threads / max(busy) / 1e6, "unit": "MS/s", "cores": threads,
            "kind": "reference" if real else "port",
            "sample": "%d independent long CFastFIR streams (one per host thread), %d passes over 2^20 samples in 2^18-sample "
                      "calls; %s" % (threads, passes, "reference pebblelib CFastFIR compiled at 131072/65537 + FFTW 3.3.4 "
                                     "(oracle/_ref/libpebbleref_long.so)" if real else "C restatement oracle/pebble_oracle.c")}


def longfir_parity(x_head, y_head):
    """The first samples stream 0 produced in the run's first step vs the reference CFastFIR compiled at 131072/65537
    (oracle/_ref/libpebbleref_long.so) — or the C restatement where oracle/_ref is absent — on the same input."""
    import numpy as np
    from oracle import ref, port
    real = ref.long_available()
    o = ref.CFastFIRLong() if real else port.CFastFIR(LF_N, LF_FIR)
    o.SetupParameters(-20e6, 20e6, 0.0, 100e6)
    want = np.zeros(len(x_head) + LF_N, dtype=np.complex128)
    k = o.ProcessData(len(x_head), np.ascontiguousarray(x_head), want)
    m = min(k, len(y_head))
    den = float(np.sqrt(np.mean(np.abs(want[:m]) ** 2)))
    err = float(np.sqrt(np.mean(np.abs(y_head[:m] - want[:m]) ** 2))) / (den if den > 0 else 1.0)
    return {"rel_rms": err, "samples": int(m), "oracle": "reference" if real else "port", "tolerance": 1e-10,
            "what": "first %d output samples of stream 0 in the run's first step vs %s on the same input" %
                    (m, "the reference CFastFIR compiled at 131072/65537 + FFTW 3.3.4" if real else "the C restatement")}


def measure(args, bench, torch, dist, dev, world, rank, steps, warmup, cpu_baseline=False, e2e=False):
    """Times the stream-sharded long filter on an already initialised process group. Returns the result dict on rank 0."""
    import numpy as np
    import pebblesdr_b200 as pb
    from pebblesdr_b200 import lib
    from pebblesdr_b200._lib import check
    from pebblesdr_b200.dist import max_over_ranks, shard_range

    n = 1 << (args.log2n if (args.log2n_set and args.workload == "longfir") else 22)          # samples per stream per step
    first, end = shard_range(LF_STREAMS, world, rank)
    S = end - first
    K, W = steps, max(warmup, 1)
    x = torch.empty(S, n, dtype=torch.complex128, device=dev)
    for s in range(S):                                          # stream `first + s`: seed 5000 + stream
        g = torch.Generator(device=dev)
        g.manual_seed(5000 + first + s)
        x[s] = torch.view_as_complex(torch.randn(n, 2, device=dev, dtype=torch.float64, generator=g) * 0.1)
    out = torch.zeros(S, n, dtype=torch.complex128, device=dev)
    op = pb.CFastFIR(LF_N, LF_FIR)
    assert op.SetupParameters(-20e6, 20e6, 0.0, 100e6) == 0
    check(lib.pebble_fastfir_set_streams(op._h, S))
    stream_ptr = torch.cuda.current_stream().cuda_stream

    def step():
        k = check(lib.pebble_fastfir_process_batch_device(op._h, n, x.data_ptr(), n, out.data_ptr(), n, stream_ptr))
        assert k == n

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    y_head = None
    npar = min(n, 1 << 20)
    for w in range(W):
        step()
        if w == 0 and rank == 0:
            y_head = out[0, :npar].cpu().numpy()                # first step of the stream: what the parity check replays
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
        step()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1), dev)
    launches = lib.pebble_launch_count() - l0
    clk = clocks.stop() if rank == 0 else None
    chk = float(torch.view_as_real(out[:, -4096:]).abs().sum().item())
    assert chk > 0 and chk == chk

    ms_e2e, m = None, min(n, 1 << 22)
    if e2e:
        # ---- end to end: stream 0 from a PAGE-LOCKED host buffer: H2D + kernels + D2H per step, device-timed ----
        hx = torch.empty(m, dtype=torch.complex128, pin_memory=True)
        hx.copy_(x[0, :m])
        hy = torch.zeros(m, dtype=torch.complex128, pin_memory=True)
        dx = torch.empty(m, dtype=torch.complex128, device=dev)
        dy = torch.empty(m, dtype=torch.complex128, device=dev)
        op2 = pb.CFastFIR(LF_N, LF_FIR)
        op2.SetupParameters(-20e6, 20e6, 0.0, 100e6)
        for it in range(2 + K):
            if it == 2:
                f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                barrier()
                f0.record()
            dx.copy_(hx, non_blocking=True)
            op2.ProcessData(m, dx, dy)
            hy.copy_(dy, non_blocking=True)
        f1.record()
        barrier()
        ms_e2e = max_over_ranks(f0.elapsed_time(f1), dev)

    if rank != 0:
        return None
    peak, peak_src = bench.measured_peak()
    fpk, fpk_src = bench.fp64_peak()
    total = LF_STREAMS * n                                   # whole job, all ranks
    per_gpu_sps = S * n * K / (ms * 1e-3)
    achieved = ALG_BYTES * per_gpu_sps / 1e9
    res = {
        "workload": "BASELINE config 5: long-filter stress, CFastFIR 131072-point overlap-save (65537 taps, +-20 MHz) over 100 MSPS "
                    "synthetic IQ, 64 concurrent streams sharded over %d GPU(s), no data-path collective" % world,
        "metric": "complex MS/s through the long-filter CFastFIR (131072-point overlap-save)",
        "value": total * K / (ms * 1e-3) / 1e6, "unit": "MS/s", "n_gpus": world, "steps": K, "warmup": W,
        "streams": LF_STREAMS, "streams_per_gpu": S, "samples_per_stream_per_step": n, "ms_per_step": ms / K,
        "scaling": "strong", "collective": "none (streams are independent; outputs stay on their GPU)",
        "l2": "inputs (%d MiB/step/GPU) larger than the 126 MB L2; no flush" % ((16 * n * S) >> 20),
        "parity": longfir_parity(x[0, :npar].cpu().numpy(), y_head) if y_head is not None else None,
        "roofline": {"hbm": {"achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                             "traffic": bench_traffic(S * n), "algorithmic_bytes_per_step": ALG_BYTES * S * n, "per": "GPU"},
                     "fp64": {"achieved": flop_per_sample(n) * per_gpu_sps / 1e12, "peak": fpk, "unit": "TFLOP/s",
                              "frac": flop_per_sample(n) * per_gpu_sps / 1e12 / fpk, "flop_per_sample": flop_per_sample(n),
                              "peak_source": fpk_src,
                              "note": "5 N log2 N convention over the transforms actually run: 524288-point super-blocks of 7 "
                                      "reference blocks + the remaining 131072-point blocks (340 flop/sample block by block)"},
                     "kernel": "lf_cols_kernel<16,fwd> + lf_rows_fused_kernel + lf_cols_kernel<16,inv> (four-step over 524288-point "
                               "super-blocks = 7 reference blocks each, scratch L2-resident; remainder on the 131072-point plan)"},
        "gpu_launches": int(launches), "clocks": clk,
    }
    if ms_e2e is not None:
        res["e2e"] = {"value": world * m * K / (ms_e2e * 1e-3) / 1e6, "unit": "MS/s", "h2d_bytes_per_step": 16 * m,
                      "d2h_bytes_per_step": 16 * m, "ms_per_step": ms_e2e / K,
                      "call": "one stream per GPU from page-locked host buffers: H2D + pebble_fastfir_process_device + D2H per step "
                              "(CUDA events)"}
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
            "metric": r["metric"], "value": r["value"], "unit": "MS/s", "n_gpus": world, "steps": r["steps"], "warmup": r["warmup"],
            "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": r["workload"], "streams": LF_STREAMS, "streams_per_gpu": r["streams_per_gpu"],
                       "samples_per_stream_per_step": r["samples_per_stream_per_step"], "l2": r["l2"],
                       "parallelism": "streams sharded over %d GPU(s), no data-path collective" % world},
            "e2e": r.get("e2e"),
            "roofline": dict(r["roofline"]["hbm"], bound="hbm", kernel=r["roofline"]["kernel"], fp64=r["roofline"]["fp64"]),
            "parity": r["parity"], "gpu_launches": r["gpu_launches"], "clocks": r["clocks"],
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = _cpu_longfir(args)
        bench.emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


def bench_traffic(samples):
    import json
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
            e = json.load(f).get("longfir_step", {})
        if int(e.get("samples_per_launch", -1)) == samples:
            return float(e["dram_bytes"])
    except Exception:
        pass
    return None
