#!/bin/bash
# One gpurun call: per-launch device times of one eager decode step (ncu launch list) and one
# `--set full` capture of the dominant kernel.  Each ncu run only after the same command exited 0
# without ncu.  Outputs -> gpurun_out/ (copy the summaries you keep into profiles/).
set -u
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
export PYTHONUNBUFFERED=1
LAYERS=${LAYERS:-4}
CMD="python bench.py --profile-step --layers $LAYERS"
$CMD > gpurun_out/profile_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch-list rc=$?"
if [ "${FULL:-1}" = "1" ]; then
  $CMD > gpurun_out/profile_plain2.log 2>&1 &&
  ncu --set full --clock-control none --import-source on --profile-from-start off \
      -k regex:${KERNEL:-paged_decode_kernel} -c ${COUNT:-2} -f -o gpurun_out/prof_${TAG:-decode} $CMD > gpurun_out/ncu_full.log 2>&1
  echo "full rc=$?"
fi
tail -3 gpurun_out/ncu_launches.log
exit 0
