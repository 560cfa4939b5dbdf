#!/bin/bash
# in-game 1080p frame after the last changes (4 lines per warp, k_scene_finish, side stream): launch list + ncu --set full
mkdir -p gpurun_out
python tools/game_frame_run.py 64 8 4 > gpurun_out/game_frame_plain2.log 2>&1 || { echo "plain failed"; tail -5 gpurun_out/game_frame_plain2.log; exit 1; }
cat gpurun_out/game_frame_plain2.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r01_launches_game_frame.csv \
  python tools/game_frame_run.py 64 8 2 > gpurun_out/ncu_gf2_list.log 2>&1; echo "list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_build_lines|k_gather_fog|k_scene_finish|k_model_splat|k_light_prep' -s 5 -c 5 \
  -o gpurun_out/prof_r01_game_frame2 python tools/game_frame_run.py 64 8 2 > gpurun_out/ncu_gf2_full.log 2>&1; echo "full rc=$?"
