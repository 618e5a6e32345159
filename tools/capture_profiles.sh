# Run on the GPU box (gpurun): bench first (must exit 0), then the ncu launch list, then one --set full capture.
set -e
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/cap_bench.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/cap_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"blend_|onesweep|geom_backward|preprocess_kernel|make_keys" -s 22 -c 11 \
    -f -o gpurun_out/prof_full python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/cap_ncu2.log 2>&1
echo captured
