#!/bin/bash
# ncu --set full captures (B200_PROFILING.md recipe) of: the append kernel on a full staging block,
# one chunk of the per-chunk kernels, the query / window-sized assignment kernels, and the
# assignment kernel at 1M rows.  Run under gpurun after the plain commands exited 0.
set -x
python tools/ncu_target.py 256 || exit 1
NCU="ncu --set full --clock-control none --import-source on -f"
$NCU -k regex:'k_bstage' --launch-count 1 -o gpurun_out/r01_stage python tools/ncu_target.py 256 > gpurun_out/ncu_stage.log 2>&1
$NCU -k regex:'k_bbook|k_gram_tc|k_gram_reduce|k_eig_jacobi|k_rotate_tc' --launch-skip 395 --launch-count 5 \
    -o gpurun_out/r01_chunk python tools/ncu_target.py 256 > gpurun_out/ncu_chunk.log 2>&1
$NCU -k regex:'k_assign|k_eig_big|k_gram_f64|k_cov_f64|k_merge_out|k_cov_out' \
    -o gpurun_out/r01_query python tools/ncu_target.py 256 > gpurun_out/ncu_query.log 2>&1
$NCU -k regex:'k_assign_tc' --launch-skip 3 --launch-count 1 \
    -o gpurun_out/r01_assign1m python tools/assign_time.py 1048576 > gpurun_out/ncu_assign1m.log 2>&1
for f in gpurun_out/ncu_stage.log gpurun_out/ncu_chunk.log gpurun_out/ncu_query.log gpurun_out/ncu_assign1m.log; do tail -n 2 $f; done
ls -la gpurun_out/*.ncu-rep
