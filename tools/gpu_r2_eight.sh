#!/bin/bash
# 8-GPU pass: multi-GPU tests, the box's D2H ceiling at N = 4 / 8 (the box has one NUMA node: CPU affinity changes nothing,
# measured once at N = 8), bench.py at N = 8 / 4
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
nproc; lscpu | grep -E "Model name|NUMA node|Socket"
timeout 900 python -m pytest tests/test_gpu_group.py tests/test_gpu_multi.py -m gpu -x -q -rs 2>&1 | tail -8 | tee gpurun_out/r2_pytest_multigpu.log
for n in 2 4 8; do
  P=$((20000 + RANDOM % 20000))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $P tools/d2h_ceiling.py 2>/dev/null | tail -1 | cut -c1-330
done
P=$((20000 + RANDOM % 20000))
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $P tools/d2h_ceiling.py --affinity 2>/dev/null | tail -1 | cut -c1-330
for n in 8 4; do
  P=$((20000 + RANDOM % 20000))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $P bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r2_bench_n$n.json 2> gpurun_out/r2_bench_n$n.err; echo "bench n$n rc=$?"
  python - $n <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/r2_bench_n{n}.json').read().strip().splitlines()[-1])
    print('N',n,'value',round(d['value']),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value']),'e2e ms',round(d['e2e']['ms_per_step'],2))
    print(' strong', {k:v for k,v in d['strong'].items() if k!='note'})
    mg=d['multi_gpu']; print(' edit ms', mg['edit_broadcast']['ms_per_frame_total'], 'gather ms', mg['gather_to_rank0']['ms'], 'ingress GB/s', mg['gather_to_rank0']['root_ingress_gb_s'])
    print(' rays', d['extra']['castray2'].get('mrays_s'), d['extra']['castray2'].get('e2e_mrays_s'), d['extra']['castray2'].get('strong'))
except Exception as e:
    print('N',n,'failed',e)
PY
done
