"""bench_operators.py — single-operator legs of bench.py (`python bench.py --workload fastfir|spectrum`):
BASELINE config 1 (CFastFIR 2048-point overlap-save band-pass, +-3 kHz, 2 MSPS synthetic IQ) and config 3
(SignalSpectrum path: 8192-point Blackman-Harris windowed FFT + dB conversion + 2-tap averaging, 20 MSPS IQ).
Same JSON contract as the chain line. Single stream per GPU ("replicas only" at N > 1).

fastfir : algorithmic bytes 32 B/sample (16 read + 16 written). In FP64 the kernel's binding resources are the
          FP64 pipe and the shared-memory data pipe, not HBM (profiles/r1_fastfir_history.md); the HBM view is reported
          because the contract asks for it, the FP64 instruction floor next to it.
spectrum: algorithmic bytes 16 B/sample read + 8 B/bin written = 24 B/sample at one frame per 8192 samples."""
import json
import os
import time

FF_ALG_BYTES = 32.0
SP_ALG_BYTES = 24.0
FF_FP64_INSTR_PER_SAMPLE = 119.3    # SASS count of fastfir4096x3_kernel: 1432 FP64 instr/thread x 256 threads / 3072 outputs
FP64_WARP_INSTR_CYCLES = 2.0        # B200: 16 FP64 lanes per SM sub-partition


def _ncu_traffic(key, n):
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")) as f:
            e = json.load(f).get(key, {})
        if int(e.get("samples_per_launch", -1)) == n:
            return float(e["dram_bytes"])
    except Exception:
        pass
    return None


def _cpu_fastfir(args):
    """Unmodified reference CFastFIR (pebblelib + FFTW, oracle/_ref) on every host thread, bounded sample."""
    import threading
    import numpy as np
    from oracle import ref, cpu_chain
    threads = cpu_chain.host_threads()
    m = 1 << 20
    rng = np.random.default_rng(3)
    src = ((rng.standard_normal(m) + 1j * rng.standard_normal(m)) * 0.1).astype(np.complex128)
    objs = []
    for _ in range(threads):
        o = ref.CFastFIR()
        o.SetupParameters(-3000.0, 3000.0, 0.0, 2e6)
        objs.append(o)
    passes = max(1, int(args.cpu_seconds * 25e6 / m))      # ~25 MS/s per core for the FFTW path
    busy = [0.0] * threads
    gate = threading.Barrier(threads)

    def work(i):
        out = np.zeros(m + 2048, dtype=np.complex128)
        gate.wait()
        t0 = time.perf_counter()
        for _p in range(passes):
            for q in range(0, m, 65536):
                objs[i].ProcessData(65536, src[q:q + 65536], out)
        busy[i] = time.perf_counter() - t0

    ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return {"value": m * passes * threads / max(busy) / 1e6, "unit": "MS/s", "cores": threads, "kind": "reference",
            "sample": "%d independent CFastFIR objects (one per host thread), %d passes over 2^20 samples in 65536-sample "
                      "calls, unmodified pebblelib CFastFIR + FFTW 3.3.4 (oracle/_ref)" % (threads, passes)}


def _cpu_spectrum(args, size):
    """The reference's own FFTfftw::fftSpectrum (pebblelib fft.cpp/fftw.cpp + FFTW 3.3.4, oracle/_ref) on every host thread;
    the C restatement only where oracle/_ref is absent."""
    import threading
    import numpy as np
    from oracle import port, ref, cpu_chain
    impl = ref if ref.available() else port
    threads = cpu_chain.host_threads()
    frames = 64
    rng = np.random.default_rng(4)
    src = ((rng.standard_normal(frames * size) + 1j * rng.standard_normal(frames * size)) * 0.1).astype(np.complex128)
    objs = []
    for _ in range(threads):
        o = impl.FFT()
        o.fftParams(size, 0, 20e6, size, 12)
        objs.append(o)
    passes = max(1, int(args.cpu_seconds * (20e6 if impl is ref else 4e6) / (frames * size)))
    busy = [0.0] * threads
    gate = threading.Barrier(threads)

    def work(i):
        out = np.zeros(size)
        gate.wait()
        t0 = time.perf_counter()
        for _p in range(passes):
            for f in range(frames):
                objs[i].fftSpectrum(src[f * size:(f + 1) * size], out, size)
        busy[i] = time.perf_counter() - t0

    ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return {"value": frames * size * passes * threads / max(busy) / 1e6, "unit": "MS/s", "cores": threads,
            "kind": "reference" if impl is ref else "port",
            "sample": "%d independent FFT objects (one per host thread), %d passes over %d frames of %d samples, %s "
                      "(window + FFT + unfold + dB + averaging)" % (threads, passes, frames, size,
                      "unmodified pebblelib FFTfftw::fftSpectrum + FFTW 3.3.4 (oracle/_ref)" if impl is ref else
                      "C restatement oracle/pebble_oracle.c")}


def _parity(kind, x_head, y_head, size=8192, dtot=256):
    """A slice of the run's own first-step output against the compiled reference (oracle/_ref) on the same input."""
    import numpy as np
    from oracle import ref, port
    impl = ref if ref.available() else port
    name = "reference" if impl is ref else "port"
    if kind == "fastfir":
        o = impl.CFastFIR()
        o.SetupParameters(-3000.0, 3000.0, 0.0, 2e6)
        want = np.zeros(len(x_head) + 2048, dtype=np.complex128)
        k = o.ProcessData(len(x_head), np.ascontiguousarray(x_head), want)
        want, got = want[:k], y_head[:k]
    elif kind == "decimator":
        o = impl.Decimator(10e6, 32768)
        o.buildDecimationChain(10e6, 30000)
        outs = []
        for q in range(0, len(x_head), 32768):
            w = np.zeros(32768, dtype=np.complex128)
            k = o.process(np.ascontiguousarray(x_head[q:q + 32768]), w, 32768)
            outs.append(w[:k].copy())
        want = np.concatenate(outs)
        got = y_head[:len(want)]
    else:
        o = impl.FFT()
        o.fftParams(size, 0, 20e6, size, 12)
        frames = len(x_head) // size
        want = np.zeros((frames, size))
        for f in range(frames):
            o.fftSpectrum(np.ascontiguousarray(x_head[f * size:(f + 1) * size]), want[f], size)
        # frame 0 averages with the reference's uninitialised previous-amplitude buffer (fft.cpp:348-397): compare from frame 1
        want, got = want[1:], y_head[1:frames]
        un = (want > -120.0) & (want < 0.0)
        return {"max_abs_db": float(np.max(np.abs(got[un] - want[un]))), "clipped_bins_equal": bool(np.array_equal(got[~un], want[~un])),
                "samples": int(want.size), "oracle": name, "tolerance_db": 1e-6,
                "what": "%d frames of the run's first step vs FFTfftw::fftSpectrum on the same input" % (frames - 1)}
    den = float(np.sqrt(np.mean(np.abs(want) ** 2)))
    err = float(np.sqrt(np.mean(np.abs(got - want) ** 2))) / (den if den > 0 else 1.0)
    return {"rel_rms": err, "bit_exact": bool(np.array_equal(got, want)), "samples": int(len(want)), "oracle": name,
            "tolerance": 0.0 if kind == "decimator" else 1e-10,
            "what": "first %d output samples of the run's first step vs the compiled reference on the same input" % len(want)}


def _cpu_decimator(args):
    """Unmodified reference Decimator (pebblelib/decimator.cpp against the scalar vDSP stub, oracle/_ref), one object per
    host thread, 32768-sample calls, bounded sample. Falls back to the C restatement (bit-equal) where _ref is absent."""
    import threading
    import numpy as np
    from oracle import ref, port, cpu_chain
    impl = ref if ref.available() else port
    threads = cpu_chain.host_threads()
    m, call = 1 << 20, 32768
    rng = np.random.default_rng(6)
    src = ((rng.standard_normal(m) + 1j * rng.standard_normal(m)) * 0.1).astype(np.complex128)
    objs = []
    for _ in range(threads):
        o = impl.Decimator(10e6, call)
        o.buildDecimationChain(10e6, 30000)
        objs.append(o)
    passes = max(1, int(args.cpu_seconds * 12e6 / m))      # ~12 MS/s per core for the portable path
    busy = [0.0] * threads
    gate = threading.Barrier(threads)

    def work(i):
        out = np.zeros(call, dtype=np.complex128)
        gate.wait()
        t0 = time.perf_counter()
        for _p in range(passes):
            for q in range(0, m, call):
                objs[i].process(src[q:q + call].copy(), out, call)
        busy[i] = time.perf_counter() - t0

    ts = [threading.Thread(target=work, args=(i,)) for i in range(threads)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return {"value": m * passes * threads / max(busy) / 1e6, "unit": "MS/s", "cores": threads,
            "kind": "reference" if impl is ref else "port",
            "sample": "%d independent Decimator objects (one per host thread), %d passes over 2^20 samples in 32768-sample calls, "
                      "%s" % (threads, passes, "unmodified pebblelib Decimator, portable (non-vDSP) arithmetic (oracle/_ref)"
                              if impl is ref else "C restatement oracle/pebble_oracle.c")}


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
    from pebblesdr_b200.dist import max_over_ranks

    fast = args.workload == "fastfir"
    deci = args.workload == "decimator"
    log2n = args.log2n if args.log2n_set else (26 if deci else 24)
    n = 1 << log2n
    K, W = args.steps, args.warmup
    g = torch.Generator(device=dev)
    g.manual_seed(777 + rank)
    x = torch.view_as_complex(torch.randn(n, 2, device=dev, dtype=torch.float64, generator=g) * 0.1).contiguous()
    size = 8192
    if deci:
        op = pb.Decimator(10e6, 32768)
        assert op.buildDecimationChain(10e6, 30000) == 39062.5     # CIC3, hb11 x16 (4 merged steps), hb15, hb27, hb59: /256
        dtot = op.totalDecimation()
        out = torch.zeros(n // dtot, dtype=torch.complex128, device=dev)
        step = lambda: op.process(x, out, n)
        alg = 16.0 + 16.0 / dtot
    elif fast:
        op = pb.CFastFIR()
        assert op.SetupParameters(-3000.0, 3000.0, 0.0, 2e6) == 0
        out = torch.zeros(n, dtype=torch.complex128, device=dev)
        step = lambda: op.ProcessData(n, x, out)
        alg = FF_ALG_BYTES
    else:
        op = pb.FFT()
        op.fftParams(size, 0, 20e6, size, 12)              # BLACKMANHARRIS, signalspectrum.cpp
        frames = n // size
        out = torch.zeros(frames, size, dtype=torch.float64, device=dev)
        flags = torch.zeros(frames, dtype=torch.int32, device=dev)
        step = lambda: op.fftSpectrumBatch(frames, size, x, out, flags)
        alg = SP_ALG_BYTES

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    npar = min(n, 1 << 20)
    y_head = None
    for w in range(max(W, 1)):
        step()
        if w == 0 and rank == 0:      # the stream's first step: what the parity check replays
            y_head = (out[:npar // size].cpu().numpy() if not (fast or deci) else out[:npar if fast else npar // dtot].cpu().numpy())
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
        step()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1), dev)
    launches = lib.pebble_launch_count() - l0
    clk = clocks.stop() if rank == 0 else None
    chk = float(out.abs().sum().item()) if not (fast or deci) else float(torch.view_as_real(out).abs().sum().item())
    assert chk > 0 and chk == chk

    # ---- end to end: PAGE-LOCKED host buffers, H2D + device call + D2H per step, CUDA events ----
    m = min(n, 1 << 24)
    hx = torch.empty(m, dtype=torch.complex128, pin_memory=True)
    hx.copy_(x[:m])
    dx = torch.empty(m, dtype=torch.complex128, device=dev)
    if deci:
        op2 = pb.Decimator(10e6, 32768)
        op2.buildDecimationChain(10e6, 30000)
        dy = torch.zeros(m // dtot, dtype=torch.complex128, device=dev)
        dev_step = lambda: op2.process(dx, dy, m)
    elif fast:
        op2 = pb.CFastFIR()
        op2.SetupParameters(-3000.0, 3000.0, 0.0, 2e6)
        dy = torch.zeros(m, dtype=torch.complex128, device=dev)
        dev_step = lambda: op2.ProcessData(m, dx, dy)
    else:
        op2 = pb.FFT()
        op2.fftParams(size, 0, 20e6, size, 12)
        dy = torch.zeros(m // size, size, dtype=torch.float64, device=dev)
        fl2 = torch.zeros(m // size, dtype=torch.int32, device=dev)
        dev_step = lambda: op2.fftSpectrumBatch(m // size, size, dx, dy, fl2)
    hy = torch.zeros(dy.shape, dtype=dy.dtype, pin_memory=True)
    d2h = hy.numel() * hy.element_size()
    for it in range(2 + K):
        if it == 2:
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            f0.record()
        dx.copy_(hx, non_blocking=True)
        dev_step()
        hy.copy_(dy, non_blocking=True)
    f1.record()
    barrier()
    ms_e2e = max_over_ranks(f0.elapsed_time(f1), dev)

    if rank == 0:
        peak, peak_src = bench.measured_peak()
        achieved = alg * n * K / (ms * 1e-3) / 1e9
        line = {
            "metric": "complex MS/s through " + ("the Decimator chain" if deci else "CFastFIR overlap-save band-pass" if fast else "the SignalSpectrum FFT path"),
            "value": world * n * K / (ms * 1e-3) / 1e6, "unit": "MS/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": ("SURVEY a18 (alternative front end of config 2): Decimator::process, 10 MSPS, protect 30 kHz: "
                                    "CIC3, hb11 (4 merged steps, keep every 16th), hb15, hb27, hb59 => /256, 39062.5 S/s" if deci else
                                    "BASELINE config 1: CFastFIR overlap-save band-pass (2048-point FFT, 1025 taps, +-3 kHz), "
                                    "1 channel, 2 MSPS synthetic complex IQ" if fast else
                                    "BASELINE config 3: SignalSpectrum path, 8192-point Blackman-Harris windowed FFT + dB + "
                                    "2-tap averaging, one frame per 8192 samples of 20 MSPS synthetic IQ"),
                       "samples_per_step": n, "l2": "inputs (%d MiB/step) larger than the 126 MB L2; no flush" % ((16 * n) >> 20),
                       "parallelism": "%d GPU(s), independent replicas" % world},
            "e2e": {"value": world * m * K / (ms_e2e * 1e-3) / 1e6, "unit": "MS/s", "h2d_bytes_per_step": 16 * m,
                    "d2h_bytes_per_step": d2h, "ms_per_step": ms_e2e / K,
                    "call": ("pebble_decimator_process_device" if deci else "pebble_fastfir_process_device" if fast else "pebble_fft_spectrum_batch_device") +
                            " between an H2D copy from and a D2H copy to page-locked host buffers, every step (CUDA events)"},
            "roofline": {"bound": "hbm", "kernel": "dec_chain_kernel" if deci else "fastfir4096x3_kernel (+ fastfir2048_kernel for the left-over blocks)" if fast
                         else "spectrum_kernel", "achieved": achieved / world, "peak": peak, "unit": "GB/s",
                         "frac": achieved / world / peak,
                         "traffic": _ncu_traffic("dec_chain_kernel" if deci else "fastfir4096x3_kernel" if fast else "spectrum_kernel", n), "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg * n},
            "parity": _parity("decimator" if deci else "fastfir" if fast else "spectrum", x[:npar].cpu().numpy(), y_head, size,
                              dtot if deci else 256) if y_head is not None else None,
            "gpu_launches": int(launches), "clocks": clk,
        }
        if fast:
            sm_mhz = (clk or {}).get("sm_mhz") or 1965.0
            floor_ms = FF_FP64_INSTR_PER_SAMPLE * n / 32.0 / (148 * 4) * FP64_WARP_INSTR_CYCLES / (sm_mhz * 1e3)
            line["roofline"]["fp64"] = {"instr_per_sample": FF_FP64_INSTR_PER_SAMPLE, "issue_floor_ms": floor_ms,
                                        "frac_of_issue_floor": floor_ms / (ms / K),
                                        "note": "FP64 issue and shared-memory wavefronts bound this operator above the HBM floor"}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = _cpu_decimator(args) if deci else _cpu_fastfir(args) if fast else _cpu_spectrum(args, size)
        bench.emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0
