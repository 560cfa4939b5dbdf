#!/bin/bash
# end-of-round evidence: full GPU test suite, smoke, both bench arms, launch lists + ncu --set full of the final kernels
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/r02_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep "^smoke" | tee gpurun_out/r02_smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/r02_bench_ref_final.json 2> gpurun_out/r02_bench_ref_final.err
timeout 1200 python bench.py > gpurun_out/r02_bench_final.json 2> gpurun_out/r02_bench_final.err; tail -2 gpurun_out/r02_bench_final.err
CMD="python bench.py --views 4096 --steps 2 --warmup 1 --no-extra --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r02_launches_views4096.csv $CMD > gpurun_out/r02_ncu_list.log 2>&1; echo "launch list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_walk_lines|k_gather_fog|k_line_setup|k_line_compact|k_view_prep' -s 5 -c 5 -o gpurun_out/r02_prof_terrain $CMD > gpurun_out/r02_ncu_full.log 2>&1; echo "terrain full rc=$?"
CMD2="python tools/game_frame_run.py 64 8 2"
$CMD2 > gpurun_out/r02_plain_frame.log 2>&1 || { echo "frame plain failed"; tail -5 gpurun_out/r02_plain_frame.log; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_launches_game_frame.csv $CMD2 > gpurun_out/r02_ncu_list2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_walk_coop|k_gather_fog|k_line_setup|k_line_compact|k_scene_finish|k_model_splat' -s 6 -c 6 -o gpurun_out/r02_prof_frame $CMD2 > gpurun_out/r02_ncu_full2.log 2>&1; echo "frame full rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:'k_castray2|^k_radiosity$|^k_castray$|k_ao_chunk' -c 4 -o gpurun_out/r02_prof_rays python bench.py --no-cpu --steps 1 --warmup 1 > gpurun_out/r02_ncu_full3.log 2>&1; echo "rays full rc=$?"
ls -la gpurun_out/*.ncu-rep
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02_bench_final.json').read().strip().splitlines()[-1])
r=json.loads(open('gpurun_out/r02_bench_ref_final.json').read().strip().splitlines()[-1])
print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value'],'frac',d['roofline']['frac'],'step frac',d['roofline']['step']['frac'],'ref',r['value'],'cpu_baseline',d['cpu_baseline']['value'])
print('kernels',d['roofline']['kernel_ms_per_step'],'numerics',d.get('numerics_ms_per_step'))
print('parity',d['parity_sample'])
ex=d['extra']
print('c1',ex['frame_800x600_title']['ms_per_frame'],ex['frame_800x600_title']['cpu_baseline']['ms_per_frame'],'c2',ex['frame_1080p']['ms_per_frame'],ex['frame_1080p']['cpu_baseline']['ms_per_frame'],'game',ex['frame_1080p_game']['ms_per_frame'], ex['frame_1080p_game']['kernel_ms'])
print('rays',{k:v for k,v in ex['castray2'].items() if k not in('cpu_baseline','workload','e2e_note')}, ex['castray2']['cpu_baseline']['value'])
print('castray',ex['castray']['ms_per_batch'],'ao',ex['ambient']['full_map_kernel_ms'],ex['ambient']['edits_per_s'])
rr=ex['radiosity']; print('rad',rr['full_map_kernel_ms'],rr['edits_per_s'],rr['roofline']['achieved'],rr['roofline']['frac_of_no_fma_peak'],rr['cpu_baseline']['value'])
print('vxl',ex['vxl']['device_load_ms'],'links',ex['links']['rebuild_ms'],ex['links']['remove_blocks_ms_per_call'],ex['links']['clusterize']['wall_ms'])
PY
