set -x
python -m pytest tests -m gpu -q > gpurun_out/gputest2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputest2.log; tail -25 gpurun_out/gputest2.log
python tools/pipelined_probe.py 32768 20 strict-fma > gpurun_out/pipe_probe20.log 2>&1; tail -9 gpurun_out/pipe_probe20.log
python tools/ncu_target.py --n 16384 --variants fp32:v4,fp32:ab4,fp32:ab2,strict-fma:auto --launches 2 > gpurun_out/ncu_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tk_kernel -o gpurun_out/prof_variants python tools/ncu_target.py --n 16384 --variants fp32:v4,fp32:ab4,fp32:ab2,strict-fma:auto --launches 1 > gpurun_out/ncu_run.log 2>&1
tail -5 gpurun_out/ncu_plain.log gpurun_out/ncu_run.log
