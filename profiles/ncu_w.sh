#!/bin/bash
# usage: profiles/ncu_w.sh <tag> <workload> <kernel regex> [skip]  -- one full ncu capture of one kernel of one bench workload
tag=$1; wl=$2; k=$3; skip=${4:-12}
out=gpurun_out/$tag; mkdir -p $out
ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -f -o $out/${k}_$wl \
    python bench.py --workload $wl --steps 10 --warmup 5 --e2e-steps 1 --no-cpu-baseline > $out/ncu_${k}_$wl.log 2>&1; echo "ncu $k $wl exit $?"
ncu -i $out/${k}_$wl.ncu-rep --page details > $out/${k}_${wl}_details.txt 2>&1
ncu -i $out/${k}_$wl.ncu-rep --page raw --csv > $out/${k}_${wl}_raw.csv 2>&1
