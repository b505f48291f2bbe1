#!/bin/bash
# development: ncu capture of the one-launch prologues inside bench.py (run through gpurun)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B="python bench.py --no-sweep --no-sp --no-cpu --steps 6 --warmup 3"
ncu --set full --clock-control none --import-source on -k regex:prepass_qkv -s 3 -c 1 -f -o gpurun_out/r2u_prof_prepass_qkv $B --workload c3_fp8 > gpurun_out/r2u_ncu_qkv.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:prepass_qk_kernel -s 3 -c 1 -f -o gpurun_out/r2u_prof_prepass_qk $B > gpurun_out/r2u_ncu_qk.log 2>&1
for r in r2u_prof_prepass_qkv r2u_prof_prepass_qk; do ncu -i gpurun_out/$r.ncu-rep --page raw --csv > gpurun_out/${r}_raw.csv 2>/dev/null; done
python tools/bench_prepass.py > gpurun_out/r2u_prepass.txt 2>&1; tail -20 gpurun_out/r2u_prepass.txt
ls -la gpurun_out | grep r2u
