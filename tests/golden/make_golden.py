"""Generates tests/golden/*.npz from the REAL reference (oracle/_ref/libpebbleref.so = unmodified pebblelib + FFTW
3.3.4 compiled from /root/reference by oracle/build_ref.sh). Run in the build container:

    python tests/golden/make_golden.py

The reference ships no golden vectors of its own; these fixtures are what pins the C oracle
(oracle/pebble_oracle.c) and, on the GPU box where /root/reference does not exist, the CUDA path. Inputs are
regenerated from tests/golden/signals.py (exact integer-hash noise), only outputs are stored."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)
from signals import iq  # noqa: E402

CASES = {
    # name: dict(kind=..., params...)
    "fastfir_cfg1": dict(kind="fastfir", setup=(-3000.0, 3000.0, 0.0, 2e6), n=2048 * 6, seed=11, fs=2e6,
                         tones=[(1e3, 0.5), (5e4, 0.5)], noise=0.1, calls=[2048] * 6),
    "fastfir_ragged": dict(kind="fastfir", setup=(300.0, 3000.0, 0.0, 78125.0), n=7000, seed=12, fs=78125.0,
                           tones=[(1500.0, 0.5), (-1500.0, 0.5)], noise=0.05, calls=[1, 1023, 1024, 17, 2048, 2887]),
    "dc_10msps": dict(kind="dc", rate=10e6, bw=15000.0, freq=1.2e6, n=32768 * 2, seed=21, tones=[(1.2015e6, 0.3), (-2e6, 0.5)],
                      noise=0.01, calls=[32768, 32768]),
    "dc_2msps": dict(kind="dc", rate=2e6, bw=15000.0, freq=-3.5e5, n=8192 * 3, seed=22, tones=[(-3.49e5, 0.4), (6e5, 0.5)],
                     noise=0.02, calls=[8192, 4096, 12288]),
    "fft_fwd_2048": dict(kind="fft", size=2048, n=2048, seed=31, fs=2e6, tones=[(1e5, 0.5)], noise=0.1),
    "spectrum_8192": dict(kind="spectrum", size=8192, frames=4, seed=32, fs=20e6, noise=0.01,
                          tones=[(-7e6, 10 ** -0.5), (-1e6, 0.1), (0.3e6, 10 ** -1.5), (4e6, 0.01), (9e6, 10 ** -2.5)]),
    # Decimator (pebblelib/decimator.cpp): 10 MSPS -> CIC3, hb11 (x16 merged), hb15, hb27, hb59; 2 MSPS -> hb11 (x4), hb15, hb19, hb31
    "dec_10msps": dict(kind="dec", rate=10e6, bw=30000, n=8192 * 4, seed=51, tones=[(2.0e3, 0.3), (-2e6, 0.5)], noise=0.01,
                       calls=[8192, 16384, 8192]),
    "dec_2msps": dict(kind="dec", rate=2e6, bw=30000, n=2048 * 4, seed=52, tones=[(-4e3, 0.4), (6e5, 0.5)], noise=0.02,
                      calls=[2048] * 4),
    "window_bh_8192": dict(kind="window", wtype=12, n=8192),
    "window_hamming_2048": dict(kind="window", wtype=6, n=2048),
    "cfir_lp": dict(kind="cfir", args=(0, 1.0, 50.0, 5000.0, 9000.0, 78125.0), n=4096, seed=41, fs=78125.0,
                    tones=[(1000.0, 0.5), (20000.0, 0.5)], noise=0.05),
}


def case_input(c):
    if c["kind"] in ("fastfir", "cfir", "fft"):
        return iq(c["n"], c["seed"], c["fs"], c["tones"], c["noise"])
    if c["kind"] in ("dc", "dec"):
        return iq(c["n"], c["seed"], c["rate"], c["tones"], c["noise"])
    if c["kind"] == "spectrum":
        x = iq(c["size"] * c["frames"], c["seed"], c["fs"], c["tones"], c["noise"])
        x[2 * c["size"] + 5] = 0.95 + 0.1j  # frame 2 trips the overload flag (fft.cpp:138)
        return x
    return None


def run_case(c, impl):
    """impl: module with the reference-shaped classes (oracle.ref or oracle.port). Returns dict of arrays."""
    x = case_input(c)
    k = c["kind"]
    if k == "fastfir":
        f = impl.CFastFIR()
        f.SetupParameters(*c["setup"])
        outs, counts, pos = [], [], 0
        for n in c["calls"]:
            o = np.zeros(n + 2048, dtype=np.complex128)
            r = f.ProcessData(n, np.ascontiguousarray(x[pos:pos + n]), o)
            outs.append(o[:r].copy()); counts.append(r); pos += n
        return dict(out=np.concatenate(outs), counts=np.array(counts))
    if k == "dc":
        d = impl.CDownConvert()
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
    if k == "dec":
        d = impl.Decimator(c["rate"], max(c["calls"]))
        rate = d.buildDecimationChain(c["rate"], c["bw"])
        outs, counts, pos = [], [], 0
        for n in c["calls"]:
            o = np.zeros(n, dtype=np.complex128)
            r = d.process(np.ascontiguousarray(x[pos:pos + n]), o, n)
            outs.append(o[:r].copy()); counts.append(r); pos += n
        return dict(out=np.concatenate(outs), counts=np.array(counts + [d.decBy2Stages()]), rate=np.array([rate]))
    if k == "fft":
        f = impl.FFT()
        f.fftParams(c["size"], 0, c["fs"], c["size"], 14)
        o = np.zeros(c["size"], dtype=np.complex128)
        f.fftForward(x, o, c["n"])
        return dict(out=o)
    if k == "spectrum":
        f = impl.FFT()
        f.fftParams(c["size"], 0, c["fs"], c["size"], 12)
        db, flags = [], []
        for i in range(c["frames"]):
            o = np.zeros(c["size"])
            flags.append(int(f.fftSpectrum(np.ascontiguousarray(x[i * c["size"]:(i + 1) * c["size"]]), o, c["size"])))
            db.append(o)
        # frame 0 is dropped: the reference's first frame averages with its uninitialised m_fftAmplitude buffer
        # (fft.cpp:348-397 reads the previous raw value before anything wrote it)
        return dict(out=np.concatenate(db[1:]), flags=np.array(flags[1:]))
    if k == "window":
        w, cg = impl.window(c["wtype"], c["n"])
        return dict(out=w, cg=np.array([cg]))
    if k == "cfir":
        f = impl.CFir()
        nt = f.InitLPFilter(*c["args"])
        o = np.zeros(c["n"], dtype=np.complex128)
        f.ProcessFilter(c["n"], x, o)
        return dict(out=o, ntaps=np.array([nt]))
    raise ValueError(k)


def main():
    from oracle import ref
    assert ref.available(), "oracle/_ref/libpebbleref.so missing: run oracle/build_ref.sh (needs /root/reference)"
    for name, c in CASES.items():
        r = run_case(c, ref)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **r)
        print(name, {k: v.shape for k, v in r.items()})


if __name__ == "__main__":
    main()
