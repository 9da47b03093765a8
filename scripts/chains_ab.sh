#!/bin/bash
# A/B of --fusion-chains on N GPUs (headline bench, Krum only): scripts/chains_ab.sh N "1 2 4"
N=${1:-2}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
for c in ${2:-1 2}; do
  timeout 240 $RUN --master-port $((29700 + c)) bench.py --gpus $N --steps 40 --warmup 5 --also "" --fusion-chains $c 2> gpurun_out/chains${c}_n$N.err | tail -1 > gpurun_out/chains${c}_n$N.json
  python - "$c" gpurun_out/chains${c}_n$N.json <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print("chains", sys.argv[1], d["value"], d["ms_per_step"], d["e2e"]["value"])
except Exception as exc:
    print("chains", sys.argv[1], "failed:", exc)
PY
done
