# gpurun payload: ncu --set full of the headline kernel (strict-fma, k = 4, 8 columns per lane), 16384^2
timeout 18 python tools/ncu_target.py --k 4 --specs strict-fma:auto:16384:1x1 --launches 1 > gpurun_out/ncu_v8_plain.log 2>&1 && \
timeout 30 ncu --set full --clock-control none --import-source on -k regex:"tk_kernel" -c 1 -o gpurun_out/prof_v8 python tools/ncu_target.py --k 4 --specs strict-fma:auto:16384:1x1 --launches 1 > gpurun_out/ncu_v8_run.log 2>&1
tail -1 gpurun_out/ncu_v8_plain.log; tail -1 gpurun_out/ncu_v8_run.log
