#!/bin/bash
# Round-2 ncu evidence (B200_PROFILING.md recipe), run under gpurun after the plain commands exited 0:
#   1. launch list of the bench's own device leg (per-launch times: compare SHARES);
#   2. --set full of one chunk's kernels at configs[1] (K5, K2, K3, K4) -- cache control off (in-pipeline traffic);
#   3. --set full of K2 (tiled) / K3 / K4 at the configs[3] shape (d=4096, l=256);
#   4. the fused assignment kernel at 1M rows; 5. one launch of the strict engine's per-row kernel.
set -x
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu --no-e2e --streams= --no-sharded --no-refsched"
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_launches_batched.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
NCU="ncu --set full --clock-control none --cache-control none --import-source on -f"
$NCU -k regex:'k_bbook|k_gram_tc|k_gram_reduce|k_eig_jacobi|k_rotate_tc' --launch-skip 395 --launch-count 5 \
    -o gpurun_out/r02_chunk python tools/ncu_target.py 256 > gpurun_out/ncu_chunk.log 2>&1
$NCU -k regex:'k_bbook|k_gram_tiled|k_gram_reduce|k_eig_jacobi|k_rotate_tc' --launch-skip 200 --launch-count 5 \
    -o gpurun_out/r02_chunk_c3 python tools/ncu_target.py 512 256 4096 10000 128 > gpurun_out/ncu_chunk_c3.log 2>&1
# assign_time.py 1048576 launches k_assign_tc 23 times per case: noise/labels-only, noise/with score, events/labels-only,
# events/with score -> launch 49 = the screening kernel on the event stream, launch 72 = the 3xTF32 kernel on it
$NCU -k regex:k_assign_tc --launch-skip 49 --launch-count 1 \
    -o gpurun_out/r02_assign_screen python tools/assign_time.py 1048576 > gpurun_out/ncu_assign_screen.log 2>&1
$NCU -k regex:k_assign_tc --launch-skip 72 --launch-count 1 \
    -o gpurun_out/r02_assign1m python tools/assign_time.py 1048576 > gpurun_out/ncu_assign1m.log 2>&1
$NCU -k regex:'k_swfd_rows' --launch-skip 0 --launch-count 4 \
    -o gpurun_out/r02_strict python tools/ncu_target.py 256 128 1024 2000 64 reference > gpurun_out/ncu_strict.log 2>&1
for f in gpurun_out/ncu_launches.log gpurun_out/ncu_chunk.log gpurun_out/ncu_chunk_c3.log gpurun_out/ncu_assign_screen.log gpurun_out/ncu_assign1m.log gpurun_out/ncu_strict.log; do tail -n 2 $f; done
ls -la gpurun_out/*.ncu-rep gpurun_out/r02_launches_batched.csv
