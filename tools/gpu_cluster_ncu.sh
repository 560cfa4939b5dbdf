#!/bin/bash
# ncu --set full of the clusterize kernels (k_cluster_keys, cub sort, k_cluster_union, k_cluster_label, scan, k_cluster_scatter)
mkdir -p gpurun_out
timeout 120 python tools/cluster_run.py > gpurun_out/cluster_plain.log 2>&1 || { tail gpurun_out/cluster_plain.log; exit 1; }
timeout 400 ncu --set full --clock-control none --import-source on -k regex:k_cluster -o gpurun_out/r01_cluster -f python tools/cluster_run.py > gpurun_out/cluster_ncu.log 2>&1
ncu -i gpurun_out/r01_cluster.ncu-rep --page raw --csv > gpurun_out/r01_ncu_full_clusterize.csv 2>/dev/null
tail -2 gpurun_out/cluster_plain.log; wc -c gpurun_out/r01_ncu_full_clusterize.csv
