#!/bin/bash
# Multi-GPU bench lines of one 8-GPU box (run with: gpurun --gpus 8 -- bash tools/scale_run.sh): config 3s / 3c strong scaling
# at 2 / 4 / 8 ranks, config 5 (training gradient set + NCCL all-reduce of the G / D payloads) and config 2 at 8.
mkdir -p gpurun_out
run() {  # config, ranks
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) \
      bench.py --config $1 --gpus $2 --steps 10 --warmup 3 2> gpurun_out/r02_scale_config$1_n$2.err | tail -1 > gpurun_out/r02_scale_config$1_n$2.json
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r02_scale_config$1_n$2.json").read())
    ar = d.get("allreduce") or {}
    print("config $1 n=$2 value %.4g e2e %.4g ms %.2f e2e_ms %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["e2e"]["ms_per_step"]),
          "allreduce ms %s bus GB/s %s" % (ar.get("ms_per_step"), ar.get("bus_gbs")) if ar else "")
except Exception as e:
    print("config $1 n=$2 FAILED", e)
PY
}
for n in 2 4 8; do run 3s $n; done
for n in 2 8; do run 3c $n; done
for n in 2 8; do run 5 $n; done
run 2 8
