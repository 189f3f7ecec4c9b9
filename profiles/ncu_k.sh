#!/bin/bash
# usage: profiles/ncu_k.sh <tag> <kernel regex>...   -- one full ncu capture per kernel of the default bench workload
tag=$1; shift
out=gpurun_out/$tag; mkdir -p $out
for k in "$@"; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 25 -c 1 -f -o $out/$k \
      python bench.py --workload 4lane_lidar_ogm_dagm_256 --steps 30 --warmup 5 --e2e-steps 1 --no-cpu-baseline > $out/ncu_$k.log 2>&1; echo "ncu $k exit $?"
  ncu -i $out/$k.ncu-rep --page details > $out/${k}_details.txt 2>&1
  ncu -i $out/$k.ncu-rep --page raw --csv > $out/${k}_raw.csv 2>&1
done
