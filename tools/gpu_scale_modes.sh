#!/bin/bash
# N-GPU bench line in both end-to-end modes
N=${1:-4}
for m in stream sync; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 5 --warmup 3 --no-sweep --no-cpu-baseline --e2e-mode $m 2> gpurun_out/bench_${N}_$m.err | grep '^{' > gpurun_out/bench_${N}gpu_$m.json
python - $N $m <<'PY'
import json, sys
n, m = sys.argv[1:3]
try:
    d = json.loads(open(f"gpurun_out/bench_{n}gpu_{m}.json").read().strip().splitlines()[-1])
    print("n", d["n_gpus"], m, "value", round(d["value"]), "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"]), "e2e ms", round(d["e2e"]["ms_per_step"], 2), d["e2e"]["mode"], "parity", d["multi_gpu_parity"])
except Exception as e:
    print("bench failed", e); print(open(f"gpurun_out/bench_{n}_{m}.err").read()[-1500:])
PY
done
