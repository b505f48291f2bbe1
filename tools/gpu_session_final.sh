#!/bin/bash
# the round's closing 1-GPU record: full GPU suite, smoke, both bench arms (run through gpurun)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
T=${1:-r2s}
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/${T}_pytest_gpu.log 2>&1; tail -4 gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -3 gpurun_out/${T}_smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/${T}_bench_reference.json 2> gpurun_out/${T}_bench_reference.err; tail -c 600 gpurun_out/${T}_bench_reference.json
timeout 900 python bench.py > gpurun_out/${T}_bench_c2.json 2> gpurun_out/${T}_bench_c2.err; tail -c 1500 gpurun_out/${T}_bench_c2.json; tail -3 gpurun_out/${T}_bench_c2.err
