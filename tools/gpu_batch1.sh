set -x
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; tail -2 gpurun_out/r2_bench.err
python bench.py --workload channelizer --steps 5 --warmup 2 > gpurun_out/r2_ch1.json 2> gpurun_out/r2_ch1.err; tail -2 gpurun_out/r2_ch1.err
python bench.py --workload fastfir --steps 10 --warmup 3 > gpurun_out/r2_ff.json 2> gpurun_out/r2_ff.err; tail -2 gpurun_out/r2_ff.err
python bench.py --workload spectrum --steps 10 --warmup 3 > gpurun_out/r2_sp.json 2> gpurun_out/r2_sp.err; tail -2 gpurun_out/r2_sp.err
python bench.py --workload decimator --steps 10 --warmup 3 > gpurun_out/r2_dec.json 2> gpurun_out/r2_dec.err; tail -2 gpurun_out/r2_dec.err
python bench.py --workload longfir --steps 5 --warmup 2 > gpurun_out/r2_lf.json 2> gpurun_out/r2_lf.err; tail -2 gpurun_out/r2_lf.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2_launches_channelizer.csv python bench.py --workload channelizer --steps 2 --warmup 1 --no-cpu-baseline > /dev/null 2> gpurun_out/ncu_ch.err
ncu --set full --clock-control none --import-source on -k regex:poly_fir_kernel -c 1 -o gpurun_out/r2_poly_fir python bench.py --workload channelizer --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>> gpurun_out/ncu_ch.err
ncu --set full --clock-control none --import-source on -k regex:poly_dft_kernel -c 1 -o gpurun_out/r2_poly_dft python bench.py --workload channelizer --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>> gpurun_out/ncu_ch.err
ncu --set full --clock-control none --import-source on -k regex:chan_demod_kernel -c 1 -o gpurun_out/r2_chan_demod python bench.py --workload channelizer --steps 1 --warmup 1 --no-cpu-baseline > /dev/null 2>> gpurun_out/ncu_ch.err
ls -la gpurun_out/*.ncu-rep
