# gpurun payload: ncu --set full captures of the flagship kernels (plain, strict, forwarding, norm finish).
# usage: gpurun -- 'bash tools/gpu_ncu_final.sh'
set -x
timeout 900 python -m pytest tests -m gpu -q --timeout=120 -x > gpurun_out/gputest17.log 2>&1; echo "rc=$?" >> gpurun_out/gputest17.log; tail -4 gpurun_out/gputest17.log
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref_final.log 2>&1; tail -1 gpurun_out/bench_ref_final.log | cut -c1-300
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.log 2>&1; tail -1 gpurun_out/bench_n1.log | cut -c1-300
timeout 300 python bench.py --steps 1000 --warmup 24 --no-e2e > gpurun_out/bench_n1_1000.log 2>&1; tail -1 gpurun_out/bench_n1_1000.log | cut -c1-300
timeout 300 python bench.py --steps 1000 --warmup 24 --grid 16384 > gpurun_out/bench_n1_16384_1000.log 2>&1; tail -1 gpurun_out/bench_n1_16384_1000.log | cut -c1-300
SPECS=fp32:v4:32768:1x1,strict-fma:auto:32768:1x1,fp32:v4:16384:2x1,strict-fma:auto:16384:2x1,fp32:v4:16384:1x1:conv,fp32:ab4:16384:1x1
timeout 300 python tools/ncu_target.py --specs $SPECS --launches 2 > gpurun_out/ncu_plain_final.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tk_kernel|norm_finish" -o gpurun_out/prof_final python tools/ncu_target.py --specs $SPECS --launches 1 > gpurun_out/ncu_run_final.log 2>&1
tail -7 gpurun_out/ncu_plain_final.log; tail -3 gpurun_out/ncu_run_final.log
