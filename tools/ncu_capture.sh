#!/bin/bash
# ncu --set full captures of the per-chunk kernels and of the query / assignment kernels
# (B200_PROFILING.md recipe).  Run under gpurun after `python tools/ncu_target.py` exited 0.
set -x
python tools/ncu_target.py 256 || exit 1
ncu --set full --clock-control none --import-source on \
    -k regex:'k_bstage|k_bbook|k_gram_tc|k_gram_reduce|k_eig_jacobi|k_rotate_tc' --launch-skip 474 --launch-count 6 \
    -f -o gpurun_out/r01_chunk python tools/ncu_target.py 256 > gpurun_out/ncu_chunk.log 2>&1
ncu --set full --clock-control none --import-source on \
    -k regex:'k_assign|k_eig_big|k_gram_f64|k_cov_f64|k_merge_out' \
    -f -o gpurun_out/r01_query python tools/ncu_target.py 256 > gpurun_out/ncu_query.log 2>&1
tail -3 gpurun_out/ncu_chunk.log gpurun_out/ncu_query.log
ls -la gpurun_out/*.ncu-rep
