#!/bin/bash
# One gpurun call of the development loop: GPU tests named in $GRT_CHECK_TESTS (default: all),
# then the bench once per environment setting in $GRT_CHECK_AB (';'-separated, e.g.
# "GRT_CORE=gather;GRT_CORE=scatter"; empty = none), printing the kernel times side by side.
#   tools/gpurun_retry.sh 2400 gpurun_out/call.log 'GRT_CHECK_AB="GRT_SW_KERNEL=staged;GRT_SW_KERNEL=thread" bash tools/gpu_check.sh'
mkdir -p gpurun_out
TAG=${GRT_CHECK_TAG:-check}
timeout 1800 python -m pytest ${GRT_CHECK_TESTS:-tests} -m gpu -x -q -s > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
grep "passed\|failed\|rror\|pytest exit" gpurun_out/${TAG}_pytest.log | tail -6
IFS=';' read -ra SETS <<< "$GRT_CHECK_AB"
for i in "${!SETS[@]}"; do
  env ${SETS[$i]} GRT_RFMIP_COLUMNS=${GRT_RFMIP_COLUMNS:-8} GRT_RFMIP_CPU_COLUMNS=0 timeout 900 \
      python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_$i.log 2>&1
  echo "[${SETS[$i]}] bench exit $?"
  grep '^{' gpurun_out/${TAG}_bench_$i.log | tail -1 > gpurun_out/${TAG}_bench_$i.json
  python - "$i" "$TAG" <<'PY'
import json, sys
d = json.loads(open(f"gpurun_out/{sys.argv[2]}_bench_{sys.argv[1]}.json").read())
print("  ms/step %.2f" % d["ms_per_step"], d["roofline"]["kernel_ms_per_step"],
      "| rfmip column", d["rfmip"]["kernel_ms_per_column"])
PY
done
