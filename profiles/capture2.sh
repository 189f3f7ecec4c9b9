#!/bin/bash
# Round-2 capture: GPU tests, plain bench (all workloads), ncu launch list, full captures of the named kernels.
# usage: profiles/capture2.sh <tag> [kernels...]      (outputs under gpurun_out/<tag>/)
tag=${1:-run}; shift
kernels=${@:-k_raster k_lidar k_control_paths}
out=gpurun_out/$tag
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -3 $out/pytest.log
python bench.py > $out/bench_all.json 2> $out/bench.err; echo "bench exit $?"
head -c 1500 $out/bench_all.json; echo
ncu --metrics gpu__time_duration.sum --clock-control none -c 160 --csv --log-file $out/launches.csv \
    python bench.py --steps 4 --warmup 3 --e2e-steps 1 --no-cpu-baseline > $out/ncu_list.log 2>&1; echo "ncu list exit $?"
for k in $kernels; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 25 -c 1 -f -o $out/$k \
      python bench.py --steps 30 --warmup 5 --e2e-steps 1 --no-cpu-baseline > $out/ncu_$k.log 2>&1; echo "ncu $k exit $?"
  ncu -i $out/$k.ncu-rep --page details > $out/${k}_details.txt 2>&1
  ncu -i $out/$k.ncu-rep --page raw --csv > $out/${k}_raw.csv 2>&1
done
nproc; nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
