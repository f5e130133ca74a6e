# gpurun payload: tests of the 8-column layout and its throughput against v4 (strict-fma, k = 4)
set -x
timeout 600 python -m pytest tests/test_gpu_layouts.py -m gpu -q --timeout=120 -x -k "v8" > gpurun_out/gputest_v8.log 2>&1; echo "rc=$?" >> gpurun_out/gputest_v8.log; tail -12 gpurun_out/gputest_v8.log
timeout 300 python tools/perf_probe.py --n 32768 --steps 240 --ks 4 --arith strict-fma --layouts v4,v8 --tag _v8 2>&1 | grep "^{" | cut -c1-220
timeout 300 python tools/perf_probe.py --n 16384 --steps 240 --ks 4 --arith strict-fma --layouts v4,v8 --tag _v8 2>&1 | grep "^{" | cut -c1-220
timeout 300 python tools/perf_probe.py --nx 8192 --ny 16384 --steps 240 --ks 4 --arith strict-fma --layouts v4,v8 --tag _v8 2>&1 | grep "^{" | cut -c1-220
