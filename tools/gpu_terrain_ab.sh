#!/bin/bash
# parity of the terrain path, then the quick bench (optionally with KEY=VALUE env overrides)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_terrain.py tests/test_gpu_renderer.py -m gpu -x -q 2>&1 | tail -3
bash tools/gpu_bench_quick.sh "$@"
