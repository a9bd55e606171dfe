"""Development aid: is the channelizer bit-reproducible? Runs every stage twice with fresh objects in one process and reports the
largest difference (0 = bit-identical). usage: python tools/probe_determinism.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pebblesdr_b200 as pb
n = 1 << 23
g = torch.Generator(device="cuda"); g.manual_seed(1)
x = torch.view_as_complex(torch.randn(n, 2, device="cuda", dtype=torch.float64, generator=g) * 0.05).contiguous()


def run(modes, dc_only=False):
    cz = pb.Channelizer(1024, 61.44e6, 15000.0)
    cz.setFilter(-7500.0, 7500.0)
    for k in range(1024):
        cz.setFrequency(k, -30.69e6 + 60e3 * k)
        cz.setDemodMode(k, modes[k])
    out = torch.zeros(1024, n >> 10, dtype=torch.complex128, device="cuda")
    (cz.downConvert if dc_only else cz.process)(x, n, out)
    torch.cuda.synchronize()
    return out


for name, modes, dc in (("downconvert", [7] * 1024, True), ("usb", [7] * 1024, False), ("am", [0] * 1024, False), ("nbfm", [2] * 1024, False)):
    a = run(modes, dc)
    junk = torch.empty(12345, device="cuda")      # shift later allocations
    b = run(modes, dc)
    d = (a - b).abs()
    print(name, "max diff", float(d.max()), "channels differing", int((d.max(dim=1).values > 0).sum()), flush=True)
    del junk
