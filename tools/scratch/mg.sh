# bench.py at N GPUs of one box: the contract's line (weak scaling) and the C5 sequence (strong scaling)
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-next --no-cpu > gpurun_out/r2y_bench_${N}gpu.json 2> gpurun_out/r2y_bench_${N}gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --workload c5 --steps 3 --warmup 1 > gpurun_out/r2y_c5_${N}gpu.json 2> gpurun_out/r2y_c5_${N}gpu.err
python - <<PY
import json
for f in ("gpurun_out/r2y_bench_${N}gpu.json", "gpurun_out/r2y_c5_${N}gpu.json"):
    d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
    print(f, "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), d["e2e"]["chunks_per_rank"], d["e2e"].get("device_choice"), d.get("results_sha256"))
PY
