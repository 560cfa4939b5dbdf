#!/bin/bash
# GPU suite, default bench (both arms), then launch list + ncu --set full of the model / light kernels of the in-game frame.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_r01g.json 2> gpurun_out/bench_r01g.err; echo "bench rc=$?"
python tools/game_frame_run.py 64 8 4 > gpurun_out/game_frame_plain.log 2>&1 || { echo "plain failed"; tail -5 gpurun_out/game_frame_plain.log; exit 1; }
cat gpurun_out/game_frame_plain.log
python tools/game_frame_run.py 512 32 4 > gpurun_out/game_frame_plain512.log 2>&1; cat gpurun_out/game_frame_plain512.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01_launches_game_frame.csv \
  python tools/game_frame_run.py 64 8 2 > gpurun_out/ncu_gf_list.log 2>&1; echo "list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_model|k_key_init|k_light|k_scene_finish' -s 7 -c 7 \
  -o gpurun_out/prof_r01_game_frame python tools/game_frame_run.py 64 8 2 > gpurun_out/ncu_gf_full.log 2>&1; echo "full rc=$?"
tail -2 gpurun_out/ncu_gf_full.log
