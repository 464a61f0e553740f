#!/bin/bash
# compute-sanitizer over the golden + tiny cases of the hot path, ONE tool per
# call (B200_PROFILING.md: several tools in one gpurun call have left a GPU unusable).
#   tools/sanitize.sh memcheck|racecheck|initcheck|synccheck [TAG]
# The driver below runs the 1-GPU path (parity + default options, k = 12 / 20 / 27,
# truncation, used hashes, the device-resident block hashes, FragmentFinder and its
# no-alloc variant) and the in-process multi-GPU host with 2 and 3 ranks on this GPU
# (position ranges, chunked exchange through peer pointers), each checked against the
# oracle.  Logs: gpurun_out/<TAG>_<tool>.log; copy the summary lines to profiles/.
TOOL=${1:-memcheck}
TAG=${2:-san}
mkdir -p gpurun_out
python tools/sanitize_driver.py > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
timeout 1700 compute-sanitizer --tool $TOOL --error-exitcode 7 --print-limit 20 \
    python tools/sanitize_driver.py > gpurun_out/${TAG}_${TOOL}.log 2>&1
echo "$TOOL rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|driver ok|Error|hazard" gpurun_out/${TAG}_${TOOL}.log | sort | uniq -c | head -20
