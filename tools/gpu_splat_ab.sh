#!/bin/bash
# parity of the model path, then the in-game frame's kernel times for the splat variants
mkdir -p gpurun_out
python -m pytest tests/test_gpu_models.py tests/test_gpu_renderer.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3
for n in "64 8" "512 32" "2048 64"; do
  echo "== default ($n)"; python tools/game_frame_run.py $n 6 | grep -E "splat|resolve|lights"
  for v in openspades_b200/lib/var/libosb200_*.so; do
    echo "== $v ($n)"; OSB200_LIB=$PWD/$v python tools/game_frame_run.py $n 6 | grep -E "splat"
  done
done
