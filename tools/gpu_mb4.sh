set -x
for v in "" mb4; do
  export H2D_VARIANT=$v
  [ -n "$v" ] && export H2D_NVCC_EXTRA="-DH2D_TK_MINBLOCKS=4 -DH2D_TK_KMIN=8"
  timeout 300 python tools/perf_probe.py --n 32768 --steps 400 --ks 8 --arith fp32 --layouts v4 --tag _mb$v 2>&1 | grep glups
  timeout 300 python tools/perf_probe.py --n 32768 --steps 200 --ks 8 --arith strict-fma --tag _mb$v 2>&1 | grep glups
done
