#!/bin/bash
# build_variants.sh name "nvcc extra flags" ... -> openspades_b200/lib/var/libosb200_<name>.so
mkdir -p openspades_b200/lib/var
while [ $# -ge 2 ]; do
  n=$1; f=$2; shift 2
  OSB200_LIB_OUT=$PWD/openspades_b200/lib/var/libosb200_$n.so OSB200_NVCC_EXTRA="$f" python -m openspades_b200.build > /tmp/build_$n.log 2>&1 &
done
wait
ls openspades_b200/lib/var/*.so
