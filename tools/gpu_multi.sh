#!/bin/bash
# development aid: bench.py at N ranks exactly as the driver launches it; prints the per-rank table
N=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "rc=$?"; tail -n 2 gpurun_out/bench_n$N.err | cut -c1-300
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["value"], d.get("strong_scaling"), "e2e", d["e2e"]["ms_per_step"], d["e2e"]["value"])
print(d.get("d2h_probe"))
print(d["per_rank"]["columns"])
for r in d["per_rank"]["rows"]: print([round(x,2) for x in r])
print(d["roofline"]["kernel"], d["roofline"]["frac"], d["clocks"])
PY
