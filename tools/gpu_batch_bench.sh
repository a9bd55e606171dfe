# Every bench line of the round on ONE box (gpurun -- 'bash tools/gpu_batch_bench.sh'); copies to keep go to profiles/ by hand.
set -x
O=gpurun_out/bench; mkdir -p $O
python bench.py --steps 10 --warmup 3 > $O/r2_bench.json 2> $O/r2_bench.err; tail -1 $O/r2_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $O/r2_bench_reference.json 2> $O/r2_bench_reference.err; tail -1 $O/r2_bench_reference.err
for w in fastfir spectrum decimator channelizer longfir; do
  python bench.py --workload $w --steps 10 --warmup 3 > $O/r2_bench_$w.json 2> $O/r2_bench_$w.err; tail -1 $O/r2_bench_$w.err
done
python tools/trace_chz.py 2>&1 | grep chz-trace > $O/r2_channelizer_timeline.txt
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"poly_|fastfir|chan_|ff_" -c 120 --csv --log-file $O/r2_launches_channelizer.csv python bench.py --workload channelizer --steps 2 --warmup 1 --no-cpu-baseline --no-sharded > /dev/null 2>&1
# the chain's own kernels only (the synthetic-input torch kernels would use up the launch budget)
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"dc_|fastfir|ff_|am_|cfir|agc|iq_|scale_" -c 200 --csv --log-file $O/r2_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sharded > /dev/null 2>&1
python -m pytest tests/test_gpu_dropin.py -q -m gpu 2>&1 | tail -2; cp gpurun_out/dropin_latency_*.json $O/ 2>/dev/null
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/bench/r2_bench*.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "unreadable", e); continue
    print(f.split("/")[-1], d.get("value"), d.get("unit"), "ms", d.get("ms_per_step"), "frac", (d.get("roofline") or {}).get("frac"),
          "e2e", (d.get("e2e") or {}).get("value"), "cpu", (d.get("cpu_baseline") or {}).get("value"), "par", json.dumps(d.get("parity"))[:90])
PY
