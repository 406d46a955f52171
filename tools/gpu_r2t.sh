cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
export QFLEX_KB_NOPERM=1
ncu --set full --clock-control none --import-source on -k "regex:tc_pair" -c 1 -f -o /tmp/ncu_r2t python tools/kernel_bench.py 24,4,4 > gpurun_out/r2t_ncu.log 2>&1
ncu -i /tmp/ncu_r2t.ncu-rep --page source --csv 2>/dev/null | gzip > gpurun_out/r2t_tc_pair_24_4_4_src.csv.gz
ncu -i /tmp/ncu_r2t.ncu-rep --page details --csv 2>/dev/null | gzip > gpurun_out/r2t_tc_pair_24_4_4_details.csv.gz
ls -la gpurun_out/r2t_*
