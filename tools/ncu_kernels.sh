#!/bin/bash
# One `ncu --set full` capture per shipped kernel (run under gpurun, one GPU).  Each program is first run plainly
# and profiled only if that run exits 0.  The .ncu-rep files stay on the box (they are tens of MB each); what comes
# back in gpurun_out/ are the per-launch CSV summaries of tools/ncu_summary.py (copied to profiles/ by hand) and the
# SASS-level hot spots of tools/ncu_hot.py.
# Usage: tools/ncu_kernels.sh [name ...]
#   names: stat binned binned_multi confmat class_stats rows partition exact exact_auroc reduce
set -u
mkdir -p gpurun_out
REP=${NCU_REPORT_DIR:-/tmp/ncu_reports}
mkdir -p "$REP"
# per-kernel DRAM traffic is merged into a copy of the tracked file; copy gpurun_out/traffic.json back by hand
[ -f gpurun_out/traffic.json ] || cp profiles/traffic.json gpurun_out/traffic.json 2>/dev/null
names=("$@")
[ ${#names[@]} -eq 0 ] && names=(stat binned binned_multi confmat class_stats rows partition exact exact_auroc)
for name in "${names[@]}"; do
  case $name in
    stat) log2n=28; pat='stat_scores_kernel' ;;
    binned) log2n=26; pat='binned_warphist_kernel' ;;
    binned_multi) log2n=26; pat='binned_multicol_kernel|binned_finalize_kernel' ;;
    confmat) log2n=24; pat='confmat_' ;;
    class_stats) log2n=24; pat='class_stats_kernel' ;;
    rows) log2n=26; pat='stat_scores_rows_kernel|class_stats_rows_kernel' ;;
    partition) log2n=25; pat='range_partition_kernel|range_offsets_kernel' ;;
    exact) log2n=24; pat='k4_' ;;
    exact_auroc) log2n=24; pat='k4_' ;;
    reduce) log2n=23; pat='stat_reduce_kernel' ;;
  esac
  python tools/prof_one.py $name $log2n > gpurun_out/prof_${name}_plain.log 2>&1 &&
  timeout ${NCU_TIMEOUT:-300} ncu --set full --clock-control none --import-source on -k "regex:$pat" -c 12 -f \
      -o "$REP/r2_${name}" python tools/prof_one.py $name $log2n > gpurun_out/prof_${name}_ncu.log 2>&1
  rc=$?
  if [ $rc -eq 0 ]; then
    python tools/ncu_summary.py "$REP/r2_${name}.ncu-rep" --out gpurun_out --traffic > /dev/null 2>&1
    python tools/ncu_hot.py "$REP/r2_${name}.ncu-rep" "" 25 > gpurun_out/r2_${name}_hot.txt 2>&1
  fi
  echo "$name rc=$rc"
done
