#!/bin/bash
# round 2 profile set: launch list + ncu --set full of the terrain kernels at the bench's launch size (4096 views), and of the
# single 1080p in-game frame (cooperative walker); ncu only after the plain run of the same command exited 0.
mkdir -p gpurun_out
CMD="python bench.py --views 4096 --steps 2 --warmup 1 --no-extra --no-cpu"
$CMD > gpurun_out/r02_plain.json 2> gpurun_out/r02_plain.err || { echo "plain run failed"; tail -5 gpurun_out/r02_plain.err; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r02_launches_views4096.csv $CMD > gpurun_out/r02_ncu_list.log 2>&1
echo "launch list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_walk_lines|k_gather_fog|k_line_setup|k_line_compact|k_view_prep' -s 5 -c 5 \
  -o gpurun_out/r02_prof_terrain $CMD > gpurun_out/r02_ncu_full.log 2>&1
echo "terrain full rc=$?"; tail -2 gpurun_out/r02_ncu_full.log
CMD2="python tools/game_frame_run.py 64 8 2"
$CMD2 > gpurun_out/r02_plain_frame.log 2>&1 || { echo "frame plain failed"; tail -5 gpurun_out/r02_plain_frame.log; exit 1; }
tail -2 gpurun_out/r02_plain_frame.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_launches_game_frame.csv $CMD2 > gpurun_out/r02_ncu_list2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_walk_coop|k_gather_fog|k_line_setup|k_line_compact|k_scene_finish|k_model_splat' -s 6 -c 6 \
  -o gpurun_out/r02_prof_frame $CMD2 > gpurun_out/r02_ncu_full2.log 2>&1
echo "frame full rc=$?"; ls -la gpurun_out/*.ncu-rep
