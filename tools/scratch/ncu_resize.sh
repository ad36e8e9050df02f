# ncu --set full of the 7 TMA resize launches of one bench step (after the plain run exited 0)
python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2v_plain.json 2> gpurun_out/r2v_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:'k_resize_tma' --launch-skip 21 --launch-count 7 \
    -o gpurun_out/r2v_resize_tma -f python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2v_ncu.log 2>&1
tail -2 gpurun_out/r2v_ncu.log
