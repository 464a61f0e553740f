#!/bin/bash
# Quick GPU check of a kernel change: the parity tests that exercise the sort, then
# a short bench line without the C5 block and the CPU legs.
# usage: tools/gpu_quick.sh TAG [pytest -k expression]
TAG=${1:-q}
KEXPR=${2:-"parity or configs or zz"}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -k "$KEXPR" > gpurun_out/${TAG}_tests.log 2>&1
echo "tests rc=$?"; tail -4 gpurun_out/${TAG}_tests.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu ${BENCH_ARGS} > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
echo "bench rc=$?"; tail -c 400 gpurun_out/${TAG}_bench.err
python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${TAG}_bench.json").read().strip().splitlines()[-1])
    print("value", d["value"], "ms", d["ms_per_step"], "kern", d.get("kernel_ms_per_step"), "e2e", d["e2e"]["value"])
    print("stage", {k: round(v, 4) for k, v in d["stage_ms"].items() if isinstance(v, float)})
    print("parity", d["parity"]["digest_equal"], d["default_mode"].get("digest_equal"), "default", d["default_mode"]["value"], "roofline", d["roofline"]["frac"])
    c5 = d.get("c5") or {}
    print("c5", {k: c5.get(k) for k in ("ms_per_step", "gbp_s", "digest_equal", "survey_model_frac", "error")}, [(r["anchor_size"], round(r["ms_per_step"], 2), r["digest_equal"]) for r in c5.get("k_sweep", [])])
except Exception as e:
    print("no bench line:", e)
PY
