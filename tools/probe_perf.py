"""Quick device-resident timing probe (development aid; bench.py is the contract)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pebblesdr_b200 as pb

def timeit(fn, iters=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

def main():
    logn = int(os.environ.get("PROBE_LOGN", "26"))
    n = 1 << logn
    x = torch.view_as_complex(torch.randn(n, 2, device="cuda", dtype=torch.float64) * 0.1).contiguous()
    # copy roofline on this box
    y = torch.empty_like(x)
    ms = timeit(lambda: y.copy_(x))
    print("copy %d MiB: %.3f ms  %.1f GB/s (r+w)" % (n * 16 >> 20, ms, 2 * n * 16 / ms / 1e6))
    for rate in (10e6, 2e6, 61.44e6):
        dc = pb.CDownConvert(); dc.SetDataRate(rate, 15000.0); dc.SetFrequency(1.2e5)
        S = dc.NumStages()
        out = torch.zeros(n >> S, dtype=torch.complex128, device="cuda")
        ms = timeit(lambda: dc.ProcessData(n, x, out))
        print("downconvert rate=%g S=%d: %.3f ms  %.1f MS/s  %.1f GB/s algorithmic (%.3f of 6547.5)" % (
            rate, S, ms, n / ms / 1e3, n * (16 + 16 / (1 << S)) / ms / 1e6, n * (16 + 16 / (1 << S)) / ms / 1e6 / 6547.5))
    m = 1 << min(logn, 24)
    ff = pb.CFastFIR(); ff.SetupParameters(-3000.0, 3000.0, 0.0, 2e6)
    out = torch.zeros(m, dtype=torch.complex128, device="cuda")
    ms = timeit(lambda: ff.ProcessData(m, x[:m], out))
    print("fastfir 2048: %.3f ms  %.1f MS/s  %.1f GB/s algorithmic" % (ms, m / ms / 1e3, m * 32 / ms / 1e6))

if __name__ == "__main__":
    main()
