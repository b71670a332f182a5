#!/usr/bin/env bash
# One GPU-box visit for profiles: plain run (must exit 0), ncu launch list of our kernels for one
# timed step, ncu --set full of the dominant sort pass and of the other heavy kernels.
# usage: gpu_ncu.sh [tag]
set -u
TAG=${1:-cur}
mkdir -p gpurun_out
OURS='extract|sort_|onesweep|unique|run_counts|append_rc|merge|dummy|dc_|source_prefix|scan_u32|fill_u32|flip_info|bucket_index|emit|fe_partition|pack_reads|suffix|rj_|target_nodes|ranged|filter_key'
CMD="python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_$TAG.json 2> gpurun_out/plain_$TAG.err || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.err; exit 1; }
cat gpurun_out/plain_$TAG.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:$OURS" -c 400 --csv \
   --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launch_$TAG.log 2>&1
python scripts/launch_list.py gpurun_out/launches_$TAG.csv > gpurun_out/launch_list_$TAG.txt 2>&1
tail -70 gpurun_out/launch_list_$TAG.txt
if [ "${2:-}" != "nofull" ]; then
NPS=${NCU_PASSES_PER_STEP:-34}
timeout 900 ncu --set full --clock-control none --import-source on -k regex:onesweep_pass -s $NPS -c 1 \
   -o gpurun_out/onesweep_full_$TAG -f $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
HEAVY=${NCU_HEAVY:-'extract_kernel|unique_bulk|dc_sinks|merge_tiles|emit_merged|sort_histogram'}
timeout 1200 ncu --set full --clock-control none --import-source on -k "regex:$HEAVY" -s ${NCU_HEAVY_SKIP:-11} -c ${NCU_HEAVY_COUNT:-11} \
   -o gpurun_out/heavy_full_$TAG -f $CMD > gpurun_out/ncu_heavy_$TAG.log 2>&1
ls -la gpurun_out/*.ncu-rep
fi
