"""Development aid: time the channelizer step for all-NBFM / all-AM / all-pass-through channel sets (which serial recurrence
bounds chan_demod_kernel?). usage: python tools/probe_chz.py [log2n]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pebblesdr_b200 as pb
log2n = int(sys.argv[1]) if len(sys.argv) > 1 else 24
n = 1 << log2n
g = torch.Generator(device="cuda"); g.manual_seed(1)
x = torch.view_as_complex(torch.randn(n, 2, device="cuda", dtype=torch.float64, generator=g) * 0.05).contiguous()
for name, modes in (("nbfm", [2] * 1024), ("am", [0] * 1024), ("usb", [7] * 1024), ("mixed", [2, 0] * 512)):
    cz = pb.Channelizer(1024, 61.44e6, 15000.0)
    cz.setFilter(-7500.0, 7500.0)
    for k in range(1024):
        cz.setFrequency(k, -30.69e6 + 60e3 * k)
        cz.setDemodMode(k, modes[k])
    out = torch.zeros(1024, n >> 10, dtype=torch.complex128, device="cuda")
    for _ in range(2):
        cz.process(x, n, out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        cz.process(x, n, out)
    e1.record(); torch.cuda.synchronize()
    print(name, "ms/step", e0.elapsed_time(e1) / 3, "front path", cz.frontPath(), flush=True)
