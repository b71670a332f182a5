#!/usr/bin/env bash
# quick GPU check: parity tests + one short bench line
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q ${PYTEST_ARGS:-} > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
tail -4 gpurun_out/tests.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench exit $?"
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench_quick.json'))
    print("ms_per_step", d["ms_per_step"], "value", d["value"], "roof", d["roofline"]["frac"], d["roofline"]["ms_per_launch"])
    print(d["roofline"]["phases_ms"])
except Exception as e:
    print("no bench", e); print(open('gpurun_out/bench_quick.err').read()[-2000:])
PY
if [ "${NOLIST:-}" = "" ]; then
OURS='extract|sort_|onesweep|unique|run_counts|append_rc|merge|dummy|dc_|source_prefix|scan_u32|fill_u32|flip_info|bucket_index|emit|target_nodes|ranged|filter_key|suffix'
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:$OURS" -c 400 --csv \
   --log-file gpurun_out/launches_quick.csv python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/ncu_launch_quick.log 2>&1
python scripts/launch_list.py gpurun_out/launches_quick.csv > gpurun_out/launch_list_quick.txt 2>&1
grep -v "x[0-9]* *0\.0[0-9][0-9] ms" gpurun_out/launch_list_quick.txt
fi
