#!/bin/bash
# Multi-GPU session on one box: the multi-rank GPU tests (real NCCL + peer memory),
# then bench.py at every N given.  usage: tools/gpu_multi.sh TAG "2 4 8" [tests]
TAG=$1; NS=$2; MODE=${3:-tests}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader > gpurun_out/${TAG}_gpus.txt 2>&1
if [ "$MODE" = "tests" ]; then
  timeout 1500 python -m pytest tests/test_dist_gpu.py -x -q > gpurun_out/${TAG}_tests.log 2>&1
  echo "dist tests rc=$?"; tail -3 gpurun_out/${TAG}_tests.log
fi
for N in $NS; do
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $((29500 + N)) bench.py --gpus $N --steps 10 --warmup 3 ${BENCH_ARGS} \
      > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
  echo "bench N=$N rc=$?"; tail -c 300 gpurun_out/${TAG}_bench_n$N.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/${TAG}_bench_n$N.json").read().strip().splitlines()[-1])
    print("N=$N value", round(d["value"], 2), "ms", round(d["ms_per_step"], 3), "kern", round(d.get("kernel_ms_per_step", 0), 3), "e2e", round(d["e2e"]["value"], 2), "parity", d["parity"]["digest_equal"], d["default_mode"].get("digest_equal"))
    print("  stage", {k: (round(v, 3) if isinstance(v, float) else v) for k, v in d["stage_ms"].items()})
    c5 = d.get("c5") or {}
    print("  c5", {k: c5.get(k) for k in ("ms_per_step", "gbp_s", "digest_equal", "kernel_ms_rank0", "error")}, c5.get("nvlink"), c5.get("default_mode"))
except Exception as e:
    print("no bench line:", e)
PY
done
