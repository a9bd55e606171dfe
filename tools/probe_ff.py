"""FastFIR timing probe over the kernel's env knobs (development aid). Each variant runs in its own process because the
knobs are read once. Prints time per 2^24 samples and the parity of the result against the generic kernel."""
import os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

def child():
    import torch
    import pebblesdr_b200 as pb
    m = 1 << 24
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    x = torch.view_as_complex(torch.randn(m, 2, device="cuda", dtype=torch.float64, generator=g) * 0.1).contiguous()
    ff = pb.CFastFIR(); ff.SetupParameters(-3000.0, 3000.0, 0.0, 2e6)
    out = torch.zeros(m, dtype=torch.complex128, device="cuda")
    ff.ProcessData(m, x, out); torch.cuda.synchronize()
    chk = out[: 1 << 20].clone()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): ff.ProcessData(m, x, out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    ref = os.environ.get("PROBE_REF")
    err = float("nan")
    if ref and os.path.exists(ref):
        r = torch.load(ref).cuda()
        err = float((chk - r).abs().pow(2).sum().sqrt() / r.abs().pow(2).sum().sqrt())
    elif ref:
        torch.save(chk.cpu(), ref)
    print("%-40s %.3f ms  %.1f GB/s algorithmic (%.3f of 6551)  rel-rms vs generic %.2e" % (
        os.environ.get("PROBE_TAG", ""), ms, m * 32 / ms / 1e6, m * 32 / ms / 1e6 / 6551, err), flush=True)

if __name__ == "__main__":
    if os.environ.get("PROBE_CHILD"):
        child()
    else:
        ref = "/tmp/ff_ref.pt"
        variants = [{"PEBBLE_FF_GENERIC": "1"}, {"PEBBLE_FF_NO3": "1"}] + [{"PEBBLE_FF_WAVES": str(w)} for w in (1, 2, 4)]
        for v in variants:
            env = dict(os.environ, PROBE_CHILD="1", PROBE_REF=ref, PROBE_TAG=" ".join("%s=%s" % kv for kv in v.items()), **v)
            subprocess.run([sys.executable, os.path.abspath(__file__)], env=env, check=False)
