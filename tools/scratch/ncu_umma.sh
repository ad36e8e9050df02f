# ncu --set full of the tcgen05 blur at the bench's launch shape (one launch, after the plain run exited 0)
python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2u_plain.json 2> gpurun_out/r2u_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:'k_blur_umma' --launch-skip 3 --launch-count 1 \
    -o gpurun_out/r2u_blur_umma -f python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2u_ncu.log 2>&1
tail -3 gpurun_out/r2u_ncu.log
ls -la gpurun_out/r2u_*
