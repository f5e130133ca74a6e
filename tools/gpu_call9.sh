set -x
timeout 900 python -m pytest tests -m gpu -q --timeout=120 -x > gpurun_out/gputest9.log 2>&1; echo "rc=$?" >> gpurun_out/gputest9.log; tail -6 gpurun_out/gputest9.log
XTR_TAG=_post timeout 300 python tools/xtr_probe.py 16384 96 2>&1 | tail -7
