# final build: GPU tests, the bench line, the reference arm, C5 (digest), then the ncu captures of the same command
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err || exit 1
python bench.py --impl reference > gpurun_out/r2x_bench_reference_arm.json 2> gpurun_out/r2x_ref.err
python bench.py --workload c5 --steps 3 --warmup 1 > gpurun_out/r2x_c5_1gpu.json 2> gpurun_out/r2x_c5.err
python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2x_plain.json 2> gpurun_out/r2x_plain.err || exit 1
ncu --set full --clock-control none --import-source on -k regex:'k_(resize|fast_group|distribute|blur|describe|stereo)' --launch-skip 39 --launch-count 13 \
    -o gpurun_out/r2x_full -f python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2x_ncu.log 2>&1
tail -2 gpurun_out/r2x_ncu.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2x_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-next --no-e2e > gpurun_out/r2x_ncu2.log 2>&1
python - <<PY
import json
for f in ("gpurun_out/r2x_bench.json", "gpurun_out/r2x_c5_1gpu.json", "gpurun_out/r2x_bench_reference_arm.json"):
    d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
    print(f, "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), d.get("results_sha256"), d.get("roofline", {}).get("frac"))
PY
