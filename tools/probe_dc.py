"""Down-converter timing probe (development aid): 10 MSPS chain, 2^26 samples, device-resident."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pebblesdr_b200 as pb
n = 1 << int(os.environ.get("PROBE_LOGN", "26"))
rate = float(os.environ.get("PROBE_RATE", "10e6"))
x = torch.view_as_complex(torch.randn(n, 2, device="cuda", dtype=torch.float64) * 0.1).contiguous()
dc = pb.CDownConvert(); dc.SetDataRate(rate, 15000.0); dc.SetFrequency(1.2e5)
S = dc.NumStages()
out = torch.zeros(n >> S, dtype=torch.complex128, device="cuda")
for _ in range(2): dc.ProcessData(n, x, out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): dc.ProcessData(n, x, out)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print("%s downconvert rate=%g S=%d: %.3f ms  %.1f GB/s algorithmic (%.3f of 6551)" % (
    os.environ.get("PROBE_TAG", ""), rate, S, ms, n * (16 + 16 / (1 << S)) / ms / 1e6, n * (16 + 16 / (1 << S)) / ms / 1e6 / 6551))
