mkdir -p gpurun_out/r2p
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --workload 4lane_lidar_ogm_dagm_256 > gpurun_out/r2p/bench2.log 2> gpurun_out/r2p/bench2.err; echo "bench2 exit $?"; tail -c 800 gpurun_out/r2p/bench2.err
python - <<PY
import json
l=[x for x in open("gpurun_out/r2p/bench2.log") if x.startswith("{")]
if l:
    d=json.loads(l[-1]); print(d["n_gpus"], d["value"], d["ms_per_step"], d["roofline"]["frac"], d["e2e"]["value"], d["e2e"]["value_incremental"], d["e2e"]["value_packed"], d["cpu_baseline"])
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2p/ref2.log 2>&1; echo "ref2 exit $?"; grep -c '"impl"' gpurun_out/r2p/ref2.log
