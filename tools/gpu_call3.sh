set -x
timeout 300 python -m pytest tests/test_gpu_layouts.py -m gpu -q --timeout=60 -k "strict or watchdog or nan" > gpurun_out/gputest3a.log 2>&1; echo "rc=$?" >> gpurun_out/gputest3a.log; tail -30 gpurun_out/gputest3a.log
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -q --timeout=60 -k "config1 or strict_mode_handles" > gpurun_out/gputest3b.log 2>&1; echo "rc=$?" >> gpurun_out/gputest3b.log; tail -30 gpurun_out/gputest3b.log
timeout 200 python tools/pipelined_probe.py 32768 20 strict-fma > gpurun_out/pipe_probe20.log 2>&1; tail -9 gpurun_out/pipe_probe20.log
for v in "" mix1 mix2; do H2D_VARIANT=$v timeout 120 python tools/perf_probe.py --n 32768 --steps 200 --ks 8 --layouts v4 --tag _mix$v > gpurun_out/probe_mix$v.log 2>&1; tail -1 gpurun_out/probe_mix$v.log; done
timeout 200 python tools/ncu_target.py --n 16384 --variants fp32:v4,fp32:ab4,strict-fma:auto --launches 2 > gpurun_out/ncu_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tk_kernel -o gpurun_out/prof_variants python tools/ncu_target.py --n 16384 --variants fp32:v4,fp32:ab4,strict-fma:auto --launches 1 > gpurun_out/ncu_run.log 2>&1
tail -5 gpurun_out/ncu_plain.log gpurun_out/ncu_run.log
