#!/bin/bash
# development: variant sweep + ncu evidence of the unified attention kernel (run through gpurun)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python tools/sweep_variants.py prod,kspre,turn1 --pv f8,f16 --d 128,64 --n 8192 --causal 0,1 --reps 30 > gpurun_out/r2n_variants.log 2>&1
B="python bench.py --no-sweep --no-sp --no-cpu"
$B --steps 20 --warmup 3 > gpurun_out/r2n_bench_short_c2.json 2> gpurun_out/r2n_bench_short_c2.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2n_launches_bench_c2.csv $B --steps 20 --warmup 3 > gpurun_out/r2n_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sage_attn_fwd -s 3 -c 1 -f -o gpurun_out/r2n_prof_attn_f16_c2 $B --steps 6 --warmup 3 > gpurun_out/r2n_ncu_f16.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sage_attn_fwd -s 3 -c 1 -f -o gpurun_out/r2n_prof_attn_f8_c3 $B --workload c3_fp8 --steps 6 --warmup 3 > gpurun_out/r2n_ncu_f8.log 2>&1
for r in r2n_prof_attn_f16_c2 r2n_prof_attn_f8_c3; do
  ncu -i gpurun_out/$r.ncu-rep --page raw --csv > gpurun_out/${r}_raw.csv 2>/dev/null
done
ls -la gpurun_out
