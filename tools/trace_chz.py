"""Development aid: print the channelizer's per-part stage timeline (PEBBLE_CHZ_TRACE, channelizer.cu) for the bench shape
(1024 channels, NBFM/AM alternating, 2^24 capture samples). usage: python tools/trace_chz.py [parts]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1:
    os.environ["PEBBLE_CHZ_PARTS"] = sys.argv[1]
import torch
import pebblesdr_b200 as pb
n = 1 << 24
g = torch.Generator(device="cuda"); g.manual_seed(1)
x = torch.view_as_complex(torch.randn(n, 2, device="cuda", dtype=torch.float64, generator=g) * 0.05).contiguous()
cz = pb.Channelizer(1024, 61.44e6, 15000.0)
cz.setFilter(-7500.0, 7500.0)
for k in range(1024):
    cz.setFrequency(k, -30.69e6 + 60e3 * k)
    cz.setDemodMode(k, {"am": 0, "nbfm": 2, "usb": 7}.get(os.environ.get("PEBBLE_TRACE_MODES", ""), [2, 0][k & 1]))
out = torch.zeros(1024, n >> 10, dtype=torch.complex128, device="cuda")
for _ in range(3):
    cz.process(x, n, out)
torch.cuda.synchronize()
os.environ["PEBBLE_CHZ_TRACE"] = "1"
cz.process(x, n, out)
torch.cuda.synchronize()
