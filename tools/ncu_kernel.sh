#!/bin/bash
# usage: tools/ncu_kernel.sh <tag> <kernel-regex> <m,n,k> [more shapes]
#   -> gpurun_out/ncu_<tag>.csv (raw page) and ncu_<tag>_src.csv (source page)
# One `ncu --set full` capture of the matching kernels tools/kernel_bench.py
# launches for the shapes.  The .ncu-rep is deleted (gpurun_out/ is capped at 64 MiB).
tag=$1; rx=$2; shift 2
QFLEX_KB_NOPERM=1 ncu --set full --clock-control none --import-source on -k "regex:$rx" -c 3 -f \
    -o /tmp/ncu_$tag python tools/kernel_bench.py "$@" > gpurun_out/ncu_$tag.log 2>&1
ncu -i /tmp/ncu_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_$tag.csv 2>/dev/null
ncu -i /tmp/ncu_$tag.ncu-rep --page source --csv > gpurun_out/ncu_${tag}_src.csv 2>/dev/null
rm -f /tmp/ncu_$tag.ncu-rep
