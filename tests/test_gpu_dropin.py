"""GPU test of the compiled C++ drop-in (include/pebble_dropin.hpp): tests/dropin/receiver_tail.cpp is the non-WFM branch of
Receiver::processIQData (application/receiver.cpp:826, 903-1007) written against classes with the reference's names, built
here by g++ against libpebblecuda.so and fed 2048-sample calls like the device plugin's consumer thread does. Its audio
must equal the reference chain (compiled reference steps, oracle/cpu_chain.py) + compiled CFractResampler; the per-call
latency it prints is the SURVEY 7.3 "2048-sample compat call" figure (recorded in BASELINE.md)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import rel_rms

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp):
    exe = os.path.join(tmp, "receiver_tail")
    lib = os.path.join(ROOT, "pebblesdr_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-Wall", "-I" + os.path.join(ROOT, "include"), "-o", exe,
                           os.path.join(ROOT, "tests", "dropin", "receiver_tail.cpp"), "-L" + lib, "-lpebblecuda", "-Wl,-rpath," + lib])
    return exe


@pytest.mark.parametrize("mode", ["am", "usb"])
def test_receiver_tail_compiled_against_dropin_classes(tmp_path, mode):
    from oracle import cpu_chain, ref_app
    sys.path.insert(0, ROOT)
    from bench import synth_numpy
    exe = _build(str(tmp_path))
    n = 2048 * 16 * 40
    x = synth_numpy(n)
    fin, fout = str(tmp_path / "in.c128"), str(tmp_path / "out.c128")
    x.tofile(fin)
    r = subprocess.run([exe, fin, fout, mode, "300"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    info = json.loads(r.stdout.strip().splitlines()[-1])
    got = np.fromfile(fout, dtype=np.complex128)
    kw = dict(demod_mode=0, lo=-5000.0, hi=5000.0) if mode == "am" else dict(demod_mode=7, lo=300.0, hi=3000.0)
    c = cpu_chain.CpuChain(**kw)
    demod = c.run(x.copy())
    rs = ref_app.CFractResampler(2048)
    rate = (78125 * 1.0) / (48000 * 1.0)
    want = np.concatenate([rs.Resample(demod[i:i + 2048], rate) for i in range(0, len(demod), 2048)])
    assert info["audio_samples"] == len(got) == len(want) > 0
    assert rel_rms(got, want) <= 1e-10, rel_rms(got, want)
    for k in ("per_operator_classes", "fused_pebble_receiver_process"):
        assert info[k]["median_us"] > 0
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, "dropin_latency_%s.json" % mode), "w") as f:
            json.dump(info, f)
    print(json.dumps(info))
