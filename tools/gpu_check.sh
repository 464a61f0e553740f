#!/bin/bash
# One GPU-box session: GPU tests, the bench line, and the ncu launch list of a
# short profile run.  usage: tools/gpu_check.sh TAG [tests|notests]
TAG=${1:-x}
MODE=${2:-tests}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader > gpurun_out/${TAG}_gpu.txt 2>&1
if [ "$MODE" = "tests" ]; then
  timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1
  echo "tests rc=$?" >> gpurun_out/${TAG}_tests.log
  tail -5 gpurun_out/${TAG}_tests.log
fi
timeout 1500 python bench.py --steps 10 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"
tail -c 600 gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
    print("value", d["value"], "ms", d["ms_per_step"], "kern", d.get("kernel_ms_per_step"), "e2e", d["e2e"]["value"])
    print("stage", d["stage_ms"])
    print("parity", d["parity"]["digest_equal"], d["default_mode"].get("digest_equal"), "roofline", d["roofline"]["frac"])
    c5 = d.get("c5") or {}
    print("c5", {k: c5.get(k) for k in ("ms_per_step", "gbp_s", "digest_equal", "survey_model_frac", "error")}, c5.get("k_sweep"))
except Exception as e:
    print("no bench line:", e)
PY
timeout 600 python bench.py --profile --no-c5 --steps 2 --warmup 3 > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/${TAG}_launches.csv python bench.py --profile --no-c5 --steps 2 --warmup 3 \
    > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu rc=$?"
