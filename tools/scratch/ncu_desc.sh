python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2w_plain.json 2> gpurun_out/r2w_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:'k_describe_tma' --launch-skip 3 --launch-count 1 \
    -o gpurun_out/r2w_describe_tma -f python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2w_ncu.log 2>&1
tail -2 gpurun_out/r2w_ncu.log
