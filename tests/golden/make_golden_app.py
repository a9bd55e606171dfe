"""Generates tests/golden/app_*.npz from the REAL application-side reference steps (oracle/_ref/libpebbleref_app.so =
unmodified application/{agc,demod/demod_am,demod/demod_nfm,...}.cpp, and libpebbleref_dec12.so = CDownConvert with
MAX_DECSTAGES 12; oracle/build_ref.sh steps 5-6). Run in the build container:

    python tests/golden/make_golden_app.py

These fixtures pin the C restatement (oracle/pebble_oracle.c) of AGC / Demod_AM / Demod_NFM / normalizeIQ and of the
10-stage 61.44 MSPS CDownConvert chain on machines without /root/reference. Inputs are regenerated from
tests/golden/signals.py; only outputs are stored."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)
from signals import iq, hash_noise  # noqa: E402

FS = 78125.0   # demodulator rate of BASELINE config 2
FRAME = 2048


def am_signal(n, seed, depth=0.5, steps=True):
    """AM-modulated carrier + noise with amplitude steps (exercises the AGC's peak-window rescan, agc.cpp:131-142)."""
    x = iq(n, seed, FS, [(1000.0, 0.3), (-2500.0, 0.1)], 0.02)
    t = np.arange(n) / FS
    x = x * (1.0 + depth * np.cos(2 * np.pi * 700.0 * t))
    if steps:
        env = np.ones(n)
        env[n // 4: n // 2] = 0.05
        env[n // 2: 5 * n // 8] = 2.0
        x = x * env
    return np.ascontiguousarray(x)


def fm_signal(n, seed, carrier=0.0, dev=5000.0, tone=1000.0):
    t = np.arange(n) / FS
    ph = 2 * np.pi * np.cumsum(dev * np.cos(2 * np.pi * tone * t)) / FS + 2 * np.pi * carrier * t
    return np.ascontiguousarray(0.3 * np.exp(1j * ph) + hash_noise(n, seed) * 0.01)


APP_CASES = {
    "app_agc_fast": dict(kind="agc", mode=1, thr=20, thr2=None, n=FRAME * 4, seed=61),
    "app_agc_med": dict(kind="agc", mode=2, thr=30, thr2=50, n=FRAME * 4, seed=62),
    "app_agc_slow": dict(kind="agc", mode=3, thr=40, thr2=None, n=FRAME * 4, seed=63),
    "app_agc_long": dict(kind="agc", mode=4, thr=25, thr2=10, n=FRAME * 4, seed=64),
    "app_agc_off": dict(kind="agc", mode=0, thr=37, thr2=12, n=FRAME * 2, seed=65),
    "app_am_filtered": dict(kind="am", bw=10000.0, n=FRAME * 2, seed=66),
    "app_nfm_pll": dict(kind="pll", carrier=0.0, n=FRAME * 2, seed=67),
    "app_nfm_pll_clamp": dict(kind="pll", carrier=19000.0, n=FRAME * 2, seed=68),   # beyond +-15 kHz: demod_nfm.cpp:242-245
    "app_nfm_fm2": dict(kind="fm2", n=FRAME * 2, seed=69),
    "app_dc_61msps": dict(kind="dc12", rate=61.44e6, bw=15000.0, freq=-30.69e6 + 517 * 60e3, n=32768 * 2, seed=70,
                          tones=[(-30.69e6 + 517 * 60e3 + 1500.0, 0.3), (5e6, 0.5)], noise=0.01, calls=[32768, 32768]),
    "app_normalize_iq": dict(kind="norm", n=512, seed=71),
}


class RefImpl:
    """Adapter over the compiled reference."""
    name = "reference"

    def __init__(self):
        from oracle import ref_app
        self.m = ref_app

    def agc(self, mode, thr):
        a = self.m.AGC(FS, FRAME)
        a.setAgcMode(mode, thr)
        return a, (lambda x: a.processBlock(x))

    def am(self, bw):
        d = self.m.Demod_AM(FS, FRAME)
        d.setBandwidth(bw)
        return d.processBlockFiltered

    def pll(self):
        return self.m.Demod_NFM(FS, FRAME).processBlockNCO

    def fm2(self):
        return self.m.Demod_NFM(FS, FRAME).processBlockFM2

    def dc12(self):
        return self.m.CDownConvert12()

    def norm(self, fmt, raw, order, ug, ng, rev):
        return self.m.normalizeIQ(fmt, raw, order, ug, ng, rev)


class PortImpl:
    """Adapter over the C restatement (oracle/pebble_oracle.c)."""
    name = "port"

    def __init__(self):
        from oracle import port
        self.m = port

    def agc(self, mode, thr):
        a = self.m.AGC(FS)
        a.setAgcMode(mode, thr)
        return a, (lambda x: a.processBlock(np.ascontiguousarray(x), len(x)))

    def am(self, bw):
        d = self.m.Demod(FS)
        d.setBandwidth(bw)
        self._keep = d
        return lambda x: d.am(np.ascontiguousarray(x), len(x))

    def pll(self):
        d = self.m.Demod(FS)
        self._keep = d
        return lambda x: d.nfm_pll(np.ascontiguousarray(x), len(x))

    def fm2(self):
        d = self.m.Demod(FS)
        self._keep = d
        return lambda x: d.fm2(np.ascontiguousarray(x), len(x))

    def dc12(self):
        return self.m.CDownConvert()

    def norm(self, fmt, raw, order, ug, ng, rev):
        return self.m.normalizeIQ(fmt, raw, order, ug, ng, rev)


def norm_inputs(n, seed):
    """Raw interleaved I,Q arrays for every normalizeIQ format, including the extreme codes."""
    u = hash_noise(n, seed)
    f = np.stack([u.real, u.imag], axis=1).reshape(-1)
    i8 = np.round(f * 127).astype(np.int8); i8[:4] = [-128, 127, 0, -1]
    u8 = (np.round(f * 127) + 128).astype(np.uint8); u8[:4] = [0, 255, 128, 127]
    i16 = np.round(f * 32767).astype(np.int16); i16[:4] = [-32768, 32767, 0, -1]
    f32 = f.astype(np.float32)
    f64 = f.astype(np.float64)
    return [i8, u8, i16, f32, f64]


def run_case(c, impl):
    k = c["kind"]
    if k == "agc":
        x = am_signal(c["n"], c["seed"])
        a, proc = impl.agc(c["mode"], c["thr"])
        outs = []
        for i in range(0, c["n"], FRAME):
            if c["thr2"] is not None and i == c["n"] // 2:
                a.setAgcThreshold(c["thr2"])     # mid-stream: keeps the delay line (agc.cpp:43-50)
            outs.append(proc(x[i:i + FRAME]))
        return dict(out=np.concatenate(outs))
    if k in ("am", "pll", "fm2"):
        if k == "am":
            x, f = am_signal(c["n"], c["seed"], steps=False), impl.am(c["bw"])
        elif k == "pll":
            x, f = fm_signal(c["n"], c["seed"], carrier=c["carrier"]), impl.pll()
        else:
            x, f = fm_signal(c["n"], c["seed"]), impl.fm2()
            # processBlockFM2's last sample is a function-static shared by all objects (demod_nfm.cpp:129): flush it to 0
            f(np.zeros(1, dtype=np.complex128))
        return dict(out=np.concatenate([f(x[i:i + FRAME]) for i in range(0, c["n"], FRAME)]))
    if k == "dc12":
        x = iq(c["n"], c["seed"], c["rate"], c["tones"], c["noise"])
        d = impl.dc12()
        rate = d.SetDataRate(c["rate"], c["bw"])
        d.SetFrequency(c["freq"])
        S = int(round(np.log2(c["rate"] / rate)))
        outs, pos = [], 0
        for n in c["calls"]:
            o = np.zeros(n >> S, dtype=np.complex128)
            r = d.ProcessData(n, np.ascontiguousarray(x[pos:pos + n]), o)
            assert r == n >> S
            outs.append(o.copy()); pos += n
        return dict(out=np.concatenate(outs), rate=np.array([rate]))
    if k == "norm":
        res = {}
        for fmt, raw in enumerate(norm_inputs(c["n"], c["seed"])):
            for order in range(4):
                for rev in (False, True):
                    res["f%d_o%d_r%d" % (fmt, order, int(rev))] = impl.norm(fmt, raw, order, 0.75, 1.0 / 3.0, rev)
        return res
    raise ValueError(k)


def main():
    from oracle import ref_app
    assert ref_app.available() and ref_app.dec12_available(), "run oracle/build_ref.sh (needs /root/reference)"
    impl = RefImpl()
    for name, c in APP_CASES.items():
        r = run_case(c, impl)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **r)
        print(name, {k: v.shape for k, v in list(r.items())[:3]})


if __name__ == "__main__":
    main()
