#!/bin/bash
# N-GPU evidence in one call: copy-bandwidth probe (alone vs together), then the bench line at N GPUs
N=${1:-8}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 tools/pcie_probe.py 2> gpurun_out/pcie_${N}.err | grep '^{' > gpurun_out/pcie_probe_${N}gpu.json
cat gpurun_out/pcie_probe_${N}gpu.json | cut -c1-600
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 5 --warmup 3 --no-sweep --no-cpu-baseline 2> gpurun_out/bench_${N}.err | grep '^{' > gpurun_out/bench_${N}gpu.json
python - $N <<'PY'
import json, sys
n = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/bench_{n}gpu.json").read().strip().splitlines()[-1])
    print("n", d["n_gpus"], "value", round(d["value"]), "ms", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"]), "e2e ms", round(d["e2e"]["ms_per_step"], 2), "parity", d["multi_gpu_parity"])
except Exception as e:
    print("bench failed", e); print(open(f"gpurun_out/bench_{n}.err").read()[-1500:])
PY
