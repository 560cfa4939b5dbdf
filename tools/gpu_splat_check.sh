#!/bin/bash
python -m pytest tests/test_gpu_models.py tests/test_gpu_renderer.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3
for n in "64 8" "512 32" "2048 64"; do echo "== $n"; python tools/game_frame_run.py $n 6 | grep -E "splat|setup"; done
