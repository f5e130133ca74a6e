set -x
timeout 900 python -m pytest tests -m gpu -q --timeout=120 -x > gpurun_out/gputest7.log 2>&1; echo "rc=$?" >> gpurun_out/gputest7.log; tail -6 gpurun_out/gputest7.log
XTR_TAG=_split timeout 300 python tools/xtr_probe.py 16384 96 > gpurun_out/xtr_split.log 2>&1; cat gpurun_out/xtr_split.log | tail -10
