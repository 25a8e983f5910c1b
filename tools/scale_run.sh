#!/bin/bash
# Multi-GPU bench lines of one 8-GPU box (gpurun --gpus 8 -- bash tools/scale_run.sh "3s:2 3s:4 3s:8 3c:2 ..."):
# config:ranks pairs; JSON lines land in gpurun_out/r02_scale_config<id>_n<ranks>.json
mkdir -p gpurun_out
for pair in ${1:-3s:2 3s:4 3s:8 3c:2}; do
  c=${pair%%:*}; n=${pair##*:}
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 200)) \
      bench.py --config $c --gpus $n --steps 10 --warmup 3 2> gpurun_out/r02_scale_config${c}_n$n.err | tail -1 > gpurun_out/r02_scale_config${c}_n$n.json
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/r02_scale_config${c}_n$n.json").read())
    ar = d.get("allreduce") or {}
    print("config $c n=$n value %.4g e2e %.4g ms %.2f e2e_ms %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["e2e"]["ms_per_step"]),
          "allreduce ms %s bus GB/s %s" % (ar.get("ms_per_step"), ar.get("bus_gbs")) if ar else "")
except Exception as e:
    print("config $c n=$n FAILED", e)
PY
done
