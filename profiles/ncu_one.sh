#!/bin/bash
# usage: profiles/ncu_one.sh <tag> <kernel regex> [skip]   -- one full ncu capture of one kernel of the default bench
tag=$1; k=$2; skip=${3:-25}
out=gpurun_out/$tag; mkdir -p $out
python bench.py --steps 30 --warmup 5 --e2e-steps 1 --no-cpu-baseline > $out/plain.json 2> $out/plain.err &&
ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -f -o $out/$k \
    python bench.py --steps 30 --warmup 5 --e2e-steps 1 --no-cpu-baseline > $out/ncu_$k.log 2>&1; echo "ncu $k exit $?"
ncu -i $out/$k.ncu-rep --page details > $out/${k}_details.txt 2>&1
ncu -i $out/$k.ncu-rep --page raw --csv > $out/${k}_raw.csv 2>&1
