#!/bin/bash
# 2-GPU pass: the group API and the 2-rank NCCL test (the driver's test box has one GPU), then bench.py --gpus 2
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader
timeout 1500 python -m pytest tests/test_gpu_group.py tests/test_gpu_multi.py tests/test_gpu_adapter.py tests/test_gpu_castray2.py -m gpu -x -q -rs 2>&1 | tail -15 | tee gpurun_out/r2_pytest_2gpu.log
P=$((20000 + RANDOM % 20000))
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $P bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench n2 rc=$?"; tail -3 gpurun_out/r2_bench_n2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_n2.json'))
print('value',round(d['value']),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value']))
print('strong', d['strong']); print('multi_gpu', json.dumps(d['multi_gpu'])[:1500])
print('rays', {k:v for k,v in d['extra']['castray2'].items() if k!='workload'})
PY
