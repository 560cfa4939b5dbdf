#!/bin/bash
# Last session of round 1: parity of the clusterize + cooperative VXL index changes, full GPU suite, smoke, both bench arms.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_links.py tests/test_gpu_vxl.py -x -q > gpurun_out/cl_pytest.log 2>&1
echo "links+vxl rc=$?" >> gpurun_out/cl_pytest.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/cl_pytest_all.log 2>&1
echo "all rc=$?" >> gpurun_out/cl_pytest_all.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/cl_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/cl_smoke.log
timeout 300 python bench.py > gpurun_out/cl_bench.json 2> gpurun_out/cl_bench.err
timeout 300 python bench.py --impl reference > gpurun_out/cl_bench_ref.json 2> gpurun_out/cl_bench_ref.err
for f in cl_pytest cl_pytest_all cl_smoke; do tail -n 3 gpurun_out/$f.log; done
python - <<'PY'
import json
d = json.load(open("gpurun_out/cl_bench.json"))
print(d["value"], d["e2e"]["value"], d["roofline"]["frac"])
print(json.dumps(d["extra"]["links"]["clusterize"]))
print(json.dumps(d["extra"]["vxl"]))
PY
tail -3 gpurun_out/cl_bench.err
