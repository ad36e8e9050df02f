# ncu --set full of one bench step at the bench's own launch shape (512 images per launch), after the plain run exited 0
python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2n_plain.json 2> gpurun_out/r2n_plain.err || exit 1
# launches per step: 7 k_resize, k_fast_group, 2 k_distribute, k_blur, k_describe, k_stereo = 13; skip the 3 warm-up steps
ncu --set full --clock-control none --import-source on -k regex:'k_(resize|fast_group|distribute|blur|describe|stereo)' --launch-skip 39 --launch-count 13 \
    -o gpurun_out/r2n_full -f python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2n_ncu.log 2>&1
tail -3 gpurun_out/r2n_ncu.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2n_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2n_ncu2.log 2>&1
ls -la gpurun_out/r2n_*
