#!/bin/bash
# One gpurun call: GPU tests, plain bench (both arms), ncu launch list, one full capture per hot kernel.
# usage: profiles/capture.sh <tag>      (outputs under gpurun_out/<tag>/)
tag=${1:-run}
skip=${2:-25}   # full captures take the (skip+1)-th launch of each kernel: agents are at the junction by then
out=gpurun_out/$tag
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest.log
tail -3 $out/pytest.log
python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
cat $out/bench.json
python bench.py --impl reference --steps 5 --warmup 1 > $out/bench_reference.json 2> $out/bench_ref.err; echo "ref exit $?"
cat $out/bench_reference.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 160 --csv --log-file $out/launches.csv \
    python bench.py --steps 4 --warmup 3 --e2e-steps 1 --no-cpu-baseline > $out/ncu_list.log 2>&1; echo "ncu list exit $?"
for k in k_raster k_lidar k_control_paths k_control_law k_events k_pack_bits; do
  sk=$skip
  case $k in k_events) sk=8;; k_pack_bits) sk=1;; esac   # launched by the events pass / the host-buffer call only
  ncu --set full --clock-control none --import-source on -k regex:$k -s $sk -c 1 -f -o $out/$k \
      python bench.py --steps 30 --warmup 5 --e2e-steps 1 --no-cpu-baseline > $out/ncu_$k.log 2>&1; echo "ncu $k exit $?"
  ncu -i $out/$k.ncu-rep --page details > $out/${k}_details.txt 2>&1
  ncu -i $out/$k.ncu-rep --page raw --csv > $out/${k}_raw.csv 2>&1
done
nproc; nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
