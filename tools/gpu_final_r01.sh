#!/bin/bash
# Round-1 closing run: GPU suite, both bench arms, then launch lists / ncu --set full of the in-game frame (new splat
# kernel) and of the VXL kernels.  Every ncu pass runs only after the same command exited 0 without ncu.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
python bench.py --impl reference > gpurun_out/bench_ref_r01h.json 2> gpurun_out/bench_ref_r01h.err; echo "ref rc=$?"
python bench.py > gpurun_out/bench_r01h.json 2> gpurun_out/bench_r01h.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_r01h.err
python tools/game_frame_run.py 64 8 4 > gpurun_out/game_frame_plain.log 2>&1 || { echo "game frame plain failed"; tail -5 gpurun_out/game_frame_plain.log; exit 1; }
cat gpurun_out/game_frame_plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01_launches_game_frame.csv \
  python tools/game_frame_run.py 64 8 2 > gpurun_out/ncu_gf_list.log 2>&1; echo "gf list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_model_splat' -s 1 -c 1 \
  -o gpurun_out/prof_r01_splat python tools/game_frame_run.py 64 8 2 > gpurun_out/ncu_gf_full.log 2>&1; echo "splat full rc=$?"
python tools/vxl_run.py > gpurun_out/vxl_plain.log 2>&1 || { echo "vxl plain failed"; tail -5 gpurun_out/vxl_plain.log; exit 1; }
cat gpurun_out/vxl_plain.log
ncu --set full --clock-control none --import-source on -k regex:'k_vxl|k_scan' -s 72 -c 26 \
  -o gpurun_out/prof_r01_vxl python tools/vxl_run.py > gpurun_out/ncu_vxl_full.log 2>&1; echo "vxl full rc=$?"
