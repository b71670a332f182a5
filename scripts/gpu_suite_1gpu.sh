#!/bin/bash
# One gpurun call on 1 GPU: bench lines of every BASELINE config, the reference arm, the ncu launch list
# and the --set full capture of the radix-pass kernels (exported to CSV on the box: the .ncu-rep stays there).
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 > $O/r02_bench_cfg2.json 2> $O/r02_bench_cfg2.err; echo "bench cfg2 rc=$?"; cut -c1-200 $O/r02_bench_cfg2.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r02_bench_reference.json 2> $O/r02_bench_reference.err; echo "reference rc=$?"; cut -c1-200 $O/r02_bench_reference.json
for w in cfg5 cfg4 cfg3; do
  timeout 700 python bench.py --workload $w --steps 2 --warmup 1 > $O/r02_bench_$w.json 2> $O/r02_bench_$w.err; echo "$w rc=$?"; cut -c1-260 $O/r02_bench_$w.json; tail -2 $O/r02_bench_$w.err
done
# launch list of the headline bench (cold-cache, serialised: shares, not absolutes)
python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > $O/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:mtg -c 400 --csv --log-file $O/r02_launches_cfg2.csv \
    python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > $O/ncu_launches.log 2>&1
echo "launch list rc=$?"
# the pass kernels: v3 stable / v5 stable / v5 unstable first pass (scripts/sort_pass_probe.py)
python scripts/sort_pass_probe.py > $O/r02_probe.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:onesweep_pass -c 24 -o /tmp/sort_pass \
    python scripts/sort_pass_probe.py > $O/ncu_probe.log 2>&1
echo "ncu probe rc=$?"
ncu -i /tmp/sort_pass.ncu-rep --page raw --csv > $O/r02_ncu_sort_pass_raw.csv 2>/dev/null
ncu -i /tmp/sort_pass.ncu-rep --page details --csv > $O/r02_ncu_sort_pass_details.csv 2>/dev/null
ncu -i /tmp/sort_pass.ncu-rep --page source --csv --launch-skip 1 --launch-count 1 > $O/r02_ncu_source_v3.csv 2>/dev/null
ncu -i /tmp/sort_pass.ncu-rep --page source --csv --launch-skip 16 --launch-count 1 > $O/r02_ncu_source_v5_unstable.csv 2>/dev/null
ls -la $O | tail -20; du -sh $O
cat $O/r02_probe.log
