"""Times Decimator.process (device-resident) on the GPU box: python tools/probe_dec.py [rate] [bw] [log2n]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pebblesdr_b200 as pb
rate = float(sys.argv[1]) if len(sys.argv) > 1 else 10e6
bw = float(sys.argv[2]) if len(sys.argv) > 2 else 30000
n = 1 << (int(sys.argv[3]) if len(sys.argv) > 3 else 26)
d = pb.Decimator(rate, 2048)
d.buildDecimationChain(rate, bw)
x = torch.view_as_complex(torch.randn(n, 2, dtype=torch.float64, device="cuda")).contiguous()
y = torch.zeros(n // d.totalDecimation(), dtype=torch.complex128, device="cuda")
for _ in range(3):
    d.process(x, y, n)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    d.process(x, y, n)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print("decimator chain", d.chain(), "n", n, "ms", ms, "GS/s", n / ms / 1e6, "GB/s (16 B/sample)", 16 * n / ms / 1e6)
