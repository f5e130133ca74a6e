set -x
XTR_ARITHS=strict-fma XTR_TAG=_dbg timeout 300 python tools/xtr_probe.py 16384 96 2>&1 | tail -4
XTR_ARITHS=strict-fma H2D_STRICT_FAST=0 XTR_TAG=_dbg0 timeout 300 python tools/xtr_probe.py 16384 96 2>&1 | tail -4
XTR_ARITHS=strict-fma,fp32 H2D_SPLIT_LAUNCH=0 H2D_MIN_WAVES=2 XTR_TAG=_dbg2 timeout 300 python tools/xtr_probe.py 16384 96 2>&1 | tail -7
