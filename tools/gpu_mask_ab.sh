#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_terrain.py tests/test_gpu_renderer.py tests/test_gpu_models.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3
for v in default openspades_b200/lib/var/libosb200_*.so; do
  echo "== $v"
  if [ "$v" = default ]; then unset OSB200_LIB; else export OSB200_LIB=$PWD/$v; fi
  python bench.py --steps 5 --warmup 3 --no-extra --no-cpu 2> /dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('value %.0f ms_per_step %.3f e2e %.0f kernels %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], {k: round(v,3) for k,v in d['roofline']['kernel_ms_per_step'].items()}))"
done
