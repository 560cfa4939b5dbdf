#!/bin/bash
python -m pytest tests/test_gpu_terrain.py tests/test_gpu_renderer.py tests/test_gpu_models.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3
for l in 32 16 8 4 2 1; do echo "== min lanes $l"; OSB200_LINE_MIN_LANES=$l python tools/game_frame_run.py 64 8 6 | grep -E "build_lines|gather"; done
