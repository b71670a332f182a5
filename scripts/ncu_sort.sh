#!/usr/bin/env bash
# usage: ncu_sort.sh VARIANT KERNEL_REGEX OUTNAME  -- one ncu --set full capture of a sort pass kernel
V=$1; K=$2; O=$3
export MTG_SORT_VARIANT=$V
ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -o gpurun_out/$O -f \
  python -c "
import ctypes,sys
sys.path.insert(0,'.')
import mtg_b200
lib=mtg_b200._lib.load()
ms=(ctypes.c_float*3)()
print(lib.mtg_debug_sort_bench(0, 200_000_000, 62, 0, 1, ms), list(ms))
" > gpurun_out/$O.log 2>&1
tail -2 gpurun_out/$O.log
