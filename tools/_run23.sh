mkdir -p gpurun_out
rm -f gpurun_out/sweep_*.jsonl
ARGS="--pv f8 --d 128 --n 2048,8192 --reps 30"
for v in prod turn0; do
  if [ $v = prod ]; then unset SAGE_B200_LIB; else export SAGE_B200_LIB=tools/_bin/libsage_$v.so; fi
  timeout 300 python tools/bench_kernel.py $ARGS --out gpurun_out/sweep_$v.jsonl > /dev/null 2>gpurun_out/sweep_$v.err; echo rc=$?
done
python - <<'P'
import json
L=lambda n:[json.loads(l) for l in open(f'gpurun_out/sweep_{n}.jsonl')]
for r in zip(L('prod'),L('turn0')): print(r[0]['pv'],r[0]['d'],r[0]['causal'],r[0]['n'],'prod/turn0',' '.join(f"{round(x['tops'])}" for x in r))
P
