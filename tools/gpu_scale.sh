#!/bin/bash
# bench.py at N GPUs under torchrun, as the driver launches it (both arms)
N=${1:-4}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --impl reference --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "ref rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
python - <<PY
import json
for f in ("gpurun_out/bench_ref_n$N.json", "gpurun_out/bench_n$N.json"):
    lines = [l for l in open(f).read().splitlines() if l.strip()]
    print(f, "stdout lines:", len(lines))
    d = json.loads(lines[-1])
    print({k: d.get(k) for k in ("impl", "value", "n_gpus", "ms_per_step")}, (d.get("e2e") or {}).get("value"))
PY
tail -3 gpurun_out/bench_n$N.err
