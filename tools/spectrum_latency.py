"""Latency of ONE spectrum frame through the host-pointer calls the display path makes at the waterfall rate
(SignalSpectrum::unprocessed, signalspectrum.cpp:63-86, and FFT::fftSpectrum, fftw.cpp:70-81), 8192 bins.
usage: python tools/spectrum_latency.py > profiles/spectrum_latency.json"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import pebblesdr_b200 as pb

N, CALLS = 8192, 2000
rng = np.random.default_rng(1)
x = np.ascontiguousarray((rng.standard_normal(N) + 1j * rng.standard_normal(N)) * 0.05)


def stats(f, calls=CALLS):
    for _ in range(50):
        f(0)
    t = []
    for i in range(calls):
        t0 = time.perf_counter_ns()
        f(i)
        t.append((time.perf_counter_ns() - t0) / 1e3)
    t = np.sort(np.array(t))
    return {"calls": calls, "median_us": round(float(t[len(t) // 2]), 2), "p99_us": round(float(t[int(0.99 * len(t))]), 2),
            "max_us": round(float(t[-1]), 2)}


f = pb.FFT(); f.fftParams(N, 0, 20e6, N, 12)
db = np.zeros(N)
res = {"fftSpectrum_host_call_8192_bins": stats(lambda i: f.fftSpectrum(x, db, N))}
# the reference's display shape: framesPerBuffer = 2048 samples zero-padded into 4096 bins (signalspectrum.cpp:21-52)
ss = pb.SignalSpectrum(2000000, 62500, 2048, 4096, 2048, 30)
x2 = np.ascontiguousarray(x[:2048])
# every call 40 ms of caller clock after the last one: the 30 frames/s gate lets each through
res["SignalSpectrum_unprocessed_host_call_2048_into_4096_bins"] = stats(lambda i: ss.unprocessed(x2, 2048, 40 * (i + 1)))
res["frame_period_ms_at_10_to_30_fps"] = [100.0, 33.3]
print(json.dumps(res))
