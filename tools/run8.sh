AR_MODES=5,6 timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 tools/ar_check.py > gpurun_out/ar8b.log 2>&1
grep -E "^cfg|all-reduce \+|AR_CHECK|FAIL|ok$" gpurun_out/ar8b.log
FW=$(python - <<'PY'
import re
t=open('gpurun_out/ar8b.log').read()
m=re.search(r"512x4096: two kernels\s+([\d.]+) us.*push-2\s+([\d.]+)", t)
print(8 if m and float(m.group(2)) < float(m.group(1)) else 4)
PY
)
echo "fused world $FW"
B200_AR_FUSED_WORLD=$FW timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/b8.json 2> gpurun_out/b8.err; echo rc=$?; tail -2 gpurun_out/b8.err
python -c "
import json;d=json.loads(open('gpurun_out/b8.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['e2e']['value'],d['roofline']['frac'],d['tp_check'],d['allreduce']); print([(c['name'],c.get('value'),c.get('allreduce')) for c in d['extra']['configs']])"
