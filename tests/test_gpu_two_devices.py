"""GPU test (needs two devices; skipped on a one-GPU box): handles created with different current devices in ONE process.
Launch configuration that depends on the device (SM count, opt-in shared-memory attributes) is kept per handle / per device
(FastFIR's function-static setup, the process-wide SM count), so the same operator must give the same result on
device 1 as on device 0 — including the 214 KB-shared-memory FastFIR kernel, whose attribute is per device."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_same_results_on_device_0_and_device_1():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import pebblesdr_b200 as pb
    rng = np.random.default_rng(3)
    n = 1 << 22
    x = np.ascontiguousarray((rng.standard_normal(n) + 1j * rng.standard_normal(n)) * 0.05)
    outs = []
    for dev in (0, 1, 0):
        with torch.cuda.device(dev):
            xd = torch.from_numpy(x).cuda()
            ff = pb.CFastFIR()
            assert ff.SetupParameters(-3000.0, 3000.0, 0.0, 2e6) == 0
            y = torch.zeros(n, dtype=torch.complex128, device="cuda")
            k = ff.ProcessData(n, xd, y)
            dc = pb.CDownConvert()
            dc.SetDataRate(10e6, 15000.0); dc.SetFrequency(1.2e6)
            z = torch.zeros(n >> dc.NumStages(), dtype=torch.complex128, device="cuda")
            assert dc.ProcessData(n, xd, z) == n >> dc.NumStages()
            rx = pb.Receiver(10e6, 2048)
            rx.setDataRate(15000.0); rx.setMixer(1.2e6); rx.setFilter(-5000.0, 5000.0); rx.setDemodMode(pb.Demod.dmAM); rx.setAgc(0, 20)
            ka, a = rx.processIQData(xd, n)
            torch.cuda.synchronize()
            outs.append((y[:k].cpu().numpy(), z.cpu().numpy(), a[:ka].cpu().numpy()))
    for i in (1, 2):
        for got, want in zip(outs[i], outs[0]):
            assert np.array_equal(got, want)
