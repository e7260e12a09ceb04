#!/bin/bash
# scaling evidence on one box with 8 GPUs: concurrent H2D at 1/2/4/8 ranks, --verify at 8, weak / strong / w16 lines
mkdir -p gpurun_out
T=${1:-r2}
run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 "$@"; }
for n in 1 2 4 8; do
  timeout 300 $( [ $n = 1 ] && echo python || echo "python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533" ) tools/h2d_concurrent.py 2>/dev/null | grep h2d_concurrent >> gpurun_out/${T}_h2d_concurrent.jsonl
done
cat gpurun_out/${T}_h2d_concurrent.jsonl
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --verify 2>/dev/null | grep verify > gpurun_out/${T}_verify_8gpu.json; cat gpurun_out/${T}_verify_8gpu.json
for n in 1 2 4 8; do
  L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533"
  [ $n = 1 ] && L=python
  timeout 600 $L bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | tail -n 1 > gpurun_out/${T}_weak_${n}gpu.json
  timeout 600 $L bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline --scaling strong --batch 256 2>/dev/null | tail -n 1 > gpurun_out/${T}_strong_${n}gpu.json
  timeout 600 $L bench.py --gpus $n --steps 10 --warmup 3 --no-cpu-baseline --case deit_base_ibert_w16 --batch 64 2>/dev/null | tail -n 1 > gpurun_out/${T}_w16_${n}gpu.json
  python - <<PY
import json
for k in ("weak","strong","w16"):
    try:
        d=json.loads(open("gpurun_out/${T}_%s_${n}gpu.json"%k).read())
        print("$n", k, "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "e2e_u8", d.get("e2e_uint8_input") and round(d["e2e_uint8_input"]["value"]))
    except Exception as ex: print("$n", k, "failed", ex)
PY
done
