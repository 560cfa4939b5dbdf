#!/bin/bash
# full GPU suite + both bench arms (defaults)
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/r2_pytest_gpu.log; tail -6 gpurun_out/r2_pytest_gpu.log
timeout 900 python bench.py --impl reference > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref rc=$?"; tail -2 gpurun_out/r2_bench_ref.err
timeout 1200 python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2_bench.err
python - <<'PY'
import json
try:
    r=json.load(open('gpurun_out/r2_bench_ref.json')); print('ref', round(r['value']), r['cpu_baseline']['cores'], 'cores', round(r['ms_per_step']), 'ms/step')
except Exception as e: print('ref failed', e)
d=json.load(open('gpurun_out/r2_bench.json'))
print('value',round(d['value']),'ms',round(d['ms_per_step'],3),'e2e',round(d['e2e']['value']), 'cpu', d['cpu_baseline'])
print('roofline', {k:(round(v,4) if isinstance(v,float) else v) for k,v in d['roofline'].items() if k in ('kernel','achieved','frac','traffic','step')})
print('parity', d['parity_sample']); print('strong', d['strong'])
e=d['extra']
for k,v in e.items():
    print(k, {kk:(round(vv,4) if isinstance(vv,float) else vv) for kk,vv in v.items() if kk!='workload' and not isinstance(vv,dict)})
    for kk,vv in v.items():
        if isinstance(vv,dict) and kk in ('cpu_baseline','roofline','parity','strong'): print('   ',kk,vv)
PY
