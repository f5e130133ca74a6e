# gpurun payload: smoke(), then ncu --set full of the headline kernel (strict-fma, k = 4) and its forwarding twin
set -x
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
SPECS=strict-fma:auto:32768:1x1,strict-fma:auto:16384:2x1,fp32:v4:32768:1x1
timeout 300 python tools/ncu_target.py --k 4 --specs strict-fma:auto:32768:1x1,strict-fma:auto:16384:2x1 --launches 2 > gpurun_out/ncu_k4_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"tk_kernel" -o gpurun_out/prof_k4 python tools/ncu_target.py --k 4 --specs strict-fma:auto:32768:1x1,strict-fma:auto:16384:2x1 --launches 1 > gpurun_out/ncu_k4_run.log 2>&1
tail -3 gpurun_out/ncu_k4_plain.log; tail -2 gpurun_out/ncu_k4_run.log
