#!/bin/bash
# Round-end verification on the GPU box (gpurun -- 'bash tools/final_check.sh'): GPU suite, smoke, both bench arms,
# the configs[3] bench line, the assignment timing.  Everything lands in gpurun_out/.
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_suite.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/gpu_suite.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cut -c1-400 gpurun_out/bench_n1.json
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?"; cut -c1-300 gpurun_out/bench_ref.json
timeout 900 python bench.py --config 3 > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench c3 rc=$?"; cut -c1-400 gpurun_out/bench_c3.json; tail -3 gpurun_out/bench_c3.err
timeout 300 python tools/assign_time.py > gpurun_out/assign_time.log 2>&1; echo "assign rc=$?"; cat gpurun_out/assign_time.log
