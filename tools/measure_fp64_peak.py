"""Runs tools/fp64_peak (built by pebblesdr_b200/build.py from tools/fp64_peak.cu) while sampling nvidia-smi clocks and
writes profiles/fp64_peak.json — the measured FP64 FMA peak every FP64-bound roofline in this repo is quoted against
(MEASURED_PEAKS.json is the driver's file and holds no FP64 figure)."""
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sample_clocks(stop, out):
    while not stop.is_set():
        try:
            r = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,clocks_throttle_reasons.active,power.draw",
                                "--format=csv,noheader,nounits", "-i", "0"], capture_output=True, text=True, timeout=5)
            f = [s.strip() for s in r.stdout.strip().split(",")]
            out.append((float(f[0]), float(f[1]), f[2], float(f[3])))
        except Exception:
            pass
        time.sleep(0.05)


def main():
    exe = os.path.join(ROOT, "tools", "fp64_peak")
    dst = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "fp64_peak.json")
    stop, samples = threading.Event(), []
    th = threading.Thread(target=sample_clocks, args=(stop, samples))
    th.start()
    time.sleep(0.2)
    r = subprocess.run([exe, "1500"], capture_output=True, text=True)
    stop.set()
    th.join()
    if r.returncode != 0:
        sys.exit("fp64_peak failed: " + r.stderr)
    res = json.loads(r.stdout)
    busy = sorted(s[0] for s in samples if s[3] > 300.0) or sorted(s[0] for s in samples)
    res["clocks"] = {"samples": len(samples), "sm_mhz_median_under_load": busy[len(busy) // 2] if busy else None,
                     "sm_max_mhz": samples[0][1] if samples else None,
                     "reasons": sorted({s[2] for s in samples}), "power_w_max": max((s[3] for s in samples), default=None)}
    res["how"] = "tools/fp64_peak.cu: dependent-free DFMA, 8 accumulators/thread, 2 CTAs x 512 threads per SM, CUDA events"
    res["nominal_tflops"] = res["sms"] * 64 * 2 * (res["clocks"]["sm_max_mhz"] or 1965.0) * 1e6 / 1e12
    with open(dst, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
