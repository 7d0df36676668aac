#!/bin/bash
# Multi-GPU scaling runs on one box: usage tools/scaling.sh <tag> <max_gpus>
T=${1:-r2s}; M=${2:-8}; O=gpurun_out; mkdir -p $O; P=29700
run() {  # config n
  P=$((P+1))
  if [ "$2" = 1 ]; then timeout 300 python bench.py --config $1 --no-cpu-baseline --no-roofline > $O/${T}_$1_g$2.json 2> $O/${T}_$1_g$2.err
  else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port $P bench.py --gpus $2 --config $1 --no-cpu-baseline --no-roofline > $O/${T}_$1_g$2.json 2> $O/${T}_$1_g$2.err; fi
  python - <<PY
import json
try:
    d=json.loads(open("$O/${T}_$1_g$2.json").read().strip().splitlines()[-1])
    print("$1 gpus=$2", round(d["value"],1), "|", round(d["e2e"]["value"],1), "frames/s", round(d["ms_per_step"],3), "ms/step", d["config"]["frames_per_rank"])
except Exception as e:
    print("$1 gpus=$2 ERR", e); print(open("$O/${T}_$1_g$2.err").read()[-800:])
PY
}
for n in 1 2 4 8; do [ $n -le $M ] && run C2 $n; done
for n in 1 2 4 8; do [ $n -le $M ] && run C3 $n; done
for n in 1 8; do [ $n -le $M ] && run C5 $n; done
