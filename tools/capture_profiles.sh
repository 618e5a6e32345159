# Run on the GPU box (gpurun): bench lines first (each must exit 0 without ncu), then the ncu launch list, then one --set full capture.
set -e
for w in config2 config5; do
  python bench.py --steps 30 --warmup 5 --no-config4 --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_arm.json 2> gpurun_out/bench_reference_arm.err
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-config4 --no-ref-cuda > gpurun_out/cap_bench.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-config4 --no-ref-cuda > gpurun_out/cap_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"blend_|onesweep|geom_backward|preprocess_kernel|make_keys|sort_hist|tile_ranges" -s 26 -c 13 \
    -f -o gpurun_out/prof_full python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-config4 --no-ref-cuda > gpurun_out/cap_ncu2.log 2>&1
echo captured
