#!/bin/bash
# Stage times of several builds of the library on the headline workload, one line each.
# usage: tools/gpu_variants.sh TAG name1 name2 ...   ("default" = libnpge_anchor.so)
TAG=$1; shift
mkdir -p gpurun_out
for v in "$@"; do
  if [ "$v" = "default" ]; then unset NPGE_ANCHOR_LIB; else export NPGE_ANCHOR_LIB=$PWD/npge_b200/libnpge_anchor_$v.so; fi
  timeout 600 python bench.py --profile --no-c5 --steps 10 --warmup 3 > gpurun_out/${TAG}_$v.json 2> gpurun_out/${TAG}_$v.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${TAG}_$v.json").read().strip().splitlines()[-1])
    s = d["stats"]
    print("%-10s step %.3f kern %.3f ok=%s hist %.3f pass %s" % ("$v", d["ms_per_step"], d["kernel_ms_per_step"], d["digest_equal"], s["ms_hist"], [round(x, 3) for x in s["ms_pass"][:5]]))
except Exception as e:
    print("$v failed:", e, open("gpurun_out/${TAG}_$v.err").read()[-300:])
PY
done
unset NPGE_ANCHOR_LIB
