set -x
timeout 300 python tools/xtr_probe.py 16384 96 > gpurun_out/xtr1.log 2>&1; cat gpurun_out/xtr1.log | tail -10
H2D_DEBUG_NOFWD=1 XTR_TAG=_nofwd timeout 300 python tools/xtr_probe.py 16384 96 > gpurun_out/xtr2.log 2>&1; cat gpurun_out/xtr2.log | tail -10
H2D_MIN_WAVES=1 XTR_TAG=_mw1 timeout 300 python tools/xtr_probe.py 16384 96 > gpurun_out/xtr3.log 2>&1; cat gpurun_out/xtr3.log | tail -10
timeout 200 python tools/ncu_target.py --n 16384 --variants fp32:v4,strict-fma:auto --launches 2 --px 2 --py 1 > gpurun_out/ncu_plain_xtr.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tk_kernel -c 4 -o gpurun_out/prof_xtr python tools/ncu_target.py --n 16384 --variants fp32:v4,strict-fma:auto --launches 1 --px 2 --py 1 > gpurun_out/ncu_run_xtr.log 2>&1
tail -3 gpurun_out/ncu_plain_xtr.log; tail -3 gpurun_out/ncu_run_xtr.log
