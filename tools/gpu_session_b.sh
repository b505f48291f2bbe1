#!/bin/bash
# development: 2-GPU parity of the sequence-parallel paths + the sp record of bench.py (run through gpurun --gpus 2)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
T=${1:-r2p}
timeout 600 python -m pytest tests/test_parallel_gpu.py -x -q -s > gpurun_out/${T}_pytest_parallel_2gpu.log 2>&1
tail -12 gpurun_out/${T}_pytest_parallel_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 --warmup 5 --no-cpu > gpurun_out/${T}_bench_2gpu.json 2> gpurun_out/${T}_bench_2gpu.err
python - <<P
import json
d=json.loads(open("gpurun_out/${T}_bench_2gpu.json").read().strip().splitlines()[-1])
print(json.dumps(d["sp"], indent=1))
P
tail -3 gpurun_out/${T}_bench_2gpu.err
