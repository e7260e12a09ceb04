#!/bin/bash
# the driver's N>1 launch line, for N = $1
N=${1:-2}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
echo "exit $?"; tail -n 5 gpurun_out/scale_n$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/scale_n$N.json").read().strip().splitlines()[-1])
print("n_gpus", d["n_gpus"], "value", round(d["value"]), "per gpu", round(d["images_per_s_per_gpu"]), "ms", round(d["ms_per_step"],3), "e2e", d["e2e"] and round(d["e2e"]["value"]))
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 1 --warmup 0 > gpurun_out/scale_ref_n$N.json 2> gpurun_out/scale_ref_n$N.err
echo "ref exit $?"; tail -n 2 gpurun_out/scale_ref_n$N.err; cut -c1-300 gpurun_out/scale_ref_n$N.json
