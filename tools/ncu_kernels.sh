#!/bin/bash
# One `ncu --set full` capture per shipped kernel (run under gpurun, one GPU).  Each program is first run plainly
# and profiled only if that run exits 0.  Reports land in gpurun_out/; tools/ncu_summary.py turns them into the
# CSV summaries kept under profiles/.
# Usage: tools/ncu_kernels.sh [name ...]   names: stat binned binned_multi confmat class_stats exact exact_auroc
set -u
mkdir -p gpurun_out
names=("$@")
[ ${#names[@]} -eq 0 ] && names=(stat binned binned_multi confmat class_stats exact exact_auroc)
for name in "${names[@]}"; do
  case $name in
    stat) log2n=28; pat='stat_scores_kernel' ;;
    binned) log2n=26; pat='binned_warphist_kernel' ;;
    binned_multi) log2n=26; pat='binned_multicol_kernel|binned_finalize_kernel' ;;
    confmat) log2n=24; pat='confmat_' ;;
    class_stats) log2n=24; pat='class_stats_kernel' ;;
    exact) log2n=24; pat='k4_' ;;
    exact_auroc) log2n=24; pat='k4_' ;;
  esac
  python tools/prof_one.py $name $log2n > gpurun_out/prof_${name}_plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k "regex:$pat" -c 24 -f \
      -o gpurun_out/r2_${name} python tools/prof_one.py $name $log2n > gpurun_out/prof_${name}_ncu.log 2>&1
  echo "$name rc=$?"
done
