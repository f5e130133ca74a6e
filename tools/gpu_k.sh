# gpurun payload: strict-fma throughput against the sweep depth k, three tile sizes
set -x
for n in 32768 16384 8192; do
timeout 300 python tools/perf_probe.py --n $n --steps 240 --ks 8,6,4,3,2 --arith strict-fma --tag _ksweep 2>&1 | grep "^{" | cut -c1-200
done
timeout 300 python tools/perf_probe.py --nx 8192 --ny 16384 --steps 240 --ks 8,4 --arith strict-fma --tag _ksweep 2>&1 | grep "^{" | cut -c1-200
timeout 300 python tools/perf_probe.py --n 32768 --steps 20 --ks 8,4 --arith strict-fma --tag _k20 2>&1 | grep "^{" | cut -c1-200
timeout 300 python tools/perf_probe.py --n 32768 --steps 240 --ks 8,4 --arith strict --tag _ksweep 2>&1 | grep "^{" | cut -c1-200
