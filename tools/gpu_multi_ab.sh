#!/bin/bash
# A/B of exchange variants at one N: tools/gpu_multi_ab.sh TAG N "ENV1" "ENV2" ...
TAG=$1; N=$2; shift; shift
mkdir -p gpurun_out
i=0
for V in "$@"; do
  i=$((i+1))
  env $V timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29600 + i)) bench.py --gpus $N --steps 10 --warmup 3 --no-c5 \
      > gpurun_out/${TAG}_$i.json 2> gpurun_out/${TAG}_$i.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${TAG}_$i.json").read().strip().splitlines()[-1])
    s = d["stage_ms"]
    print("[$V] N=$N value", round(d["value"], 2), "ms", round(d["ms_per_step"], 3), "kern", round(d["kernel_ms_per_step"], 3), "pass_a+exposed", round(s["pass_a"], 3), "parity", d["parity"]["digest_equal"], "e2e", round(d["e2e"]["value"], 2))
except Exception as e:
    print("[$V] failed:", e, open("gpurun_out/${TAG}_$i.err").read()[-400:])
PY
done
