set -x
timeout 1500 python -m pytest tests -m gpu -q --timeout=120 -x > gpurun_out/gputest4.log 2>&1; echo "rc=$?" >> gpurun_out/gputest4.log; tail -12 gpurun_out/gputest4.log
timeout 200 python tools/perf_probe.py --n 16384 --steps 96 --ks 8,6,4 --arith strict-fma --tag _fast2 > gpurun_out/probe_strict_fast2.log 2>&1; tail -3 gpurun_out/probe_strict_fast2.log
timeout 200 python tools/perf_probe.py --n 32768 --steps 96 --ks 8 --arith strict-fma --tag _fast2 > gpurun_out/probe_strict_fast2b.log 2>&1; tail -1 gpurun_out/probe_strict_fast2b.log
for v in mix1 mix2; do H2D_VARIANT=$v timeout 120 python tools/perf_probe.py --n 32768 --steps 200 --ks 8 --layouts v4 --tag _$v > gpurun_out/probe_$v.log 2>&1; tail -1 gpurun_out/probe_$v.log; done
