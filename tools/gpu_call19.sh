set -x
timeout 200 python tools/perf_probe.py --n 32768 --steps 120 --ks 8,6,5,4 --arith strict-fma --tag _base 2>&1 | grep "^{" | cut -c1-200
H2D_VARIANT=cf2f timeout 200 python tools/perf_probe.py --n 32768 --steps 120 --ks 8,6,5,4 --arith strict-fma --tag _cf2f 2>&1 | grep "^{" | cut -c1-200
