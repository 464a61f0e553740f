#!/bin/bash
# One `ncu --set full` capture of the four large kernels of a step (pass A, pass B,
# bucket sort, emit: the first launch of each in a short profile run), after the same
# command has exited 0 without ncu.  usage: tools/gpu_ncu_full.sh TAG
TAG=${1:-x}
mkdir -p gpurun_out
timeout 600 python bench.py --profile --no-c5 --steps 2 --warmup 3 > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on \
    -k regex:'pass_a_kernel|pass_b_kernel|bucket_sort_kernel|bucket_emit_direct_kernel' -c 4 \
    -o gpurun_out/${TAG}_full -f python bench.py --profile --no-c5 --steps 2 --warmup 3 \
    > gpurun_out/${TAG}_ncu_full.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2>/dev/null
ls -la gpurun_out/${TAG}_full*
