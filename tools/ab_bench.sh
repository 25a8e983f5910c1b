#!/bin/bash
# Developer tool: bench every lib variant given on the command line (names as built by tools/ab_build.py)
for v in "$@"; do
  NFI_B200_LIB=$PWD/nerf-from-image_b200/lib/libnfi_b200_$v.so timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', 'step', round(d['ms_per_step'],2), 'fwd', round(d['roofline']['field_fwd_launch_ms'],3), 'bwd', round(d['roofline']['launch_ms'],3))"
done
