#!/bin/bash
# One gpurun --gpus 8 call: multi-process parity WITHOUT the exchange verifier (plain timing; current and
# round-1 buffer handling), then the bench lines of every multi-GPU BASELINE config.
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
DIST_CHECK_VERIFY=0 DIST_CHECK_VARIANTS=1,0 DIST_CHECK_REPEATS=5 timeout 400 $TR --master-port 29521 tests/dist_check.py > $O/r02_dist_check_n8_plain.log 2>&1
echo "dist_check plain rc=$?"; grep -c " OK " $O/r02_dist_check_n8_plain.log; grep -E "MISMATCH|summary" $O/r02_dist_check_n8_plain.log | head
MTG_SYMM_LEGACY=1 DIST_CHECK_VERIFY=0 DIST_CHECK_VARIANTS=1 DIST_CHECK_REPEATS=5 timeout 400 $TR --master-port 29522 tests/dist_check.py > $O/r02_dist_check_n8_plain_legacy.log 2>&1
echo "dist_check plain legacy rc=$?"; grep -c " OK " $O/r02_dist_check_n8_plain_legacy.log; grep -E "MISMATCH|summary" $O/r02_dist_check_n8_plain_legacy.log | head
timeout 400 $TR --master-port 29523 bench.py --gpus 8 --steps 5 --warmup 3 > $O/r02_bench_cfg2_n8.json 2> $O/r02_bench_cfg2_n8.err
echo "bench cfg2 rc=$?"; cut -c1-200 $O/r02_bench_cfg2_n8.json
p=29524
for w in cfg3 cfg4 cfg5; do
  timeout 500 $TR --master-port $p bench.py --gpus 8 --workload $w --steps 3 --warmup 2 --no-parity-check > $O/r02_bench_${w}_n8.json 2> $O/r02_bench_${w}_n8.err
  echo "bench $w rc=$?"; cut -c1-260 $O/r02_bench_${w}_n8.json; tail -2 $O/r02_bench_${w}_n8.err
  p=$((p+1))
done
du -sh $O
