set -x
python -m pytest tests -m gpu -x -q > gpurun_out/gputest1.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gputest1.log; tail -15 gpurun_out/gputest1.log
python tools/perf_probe.py --n 32768 --steps 200 --ks 8 --layouts v4,ab2,ab4 --tag _lay > gpurun_out/probe_lay.log 2>&1; tail -4 gpurun_out/probe_lay.log
python tools/perf_probe.py --n 16384 --steps 96 --ks 8,4 --arith strict-fma --tag _fast > gpurun_out/probe_strict_fast.log 2>&1; tail -3 gpurun_out/probe_strict_fast.log
H2D_STRICT_FAST=0 python tools/perf_probe.py --n 16384 --steps 96 --ks 8 --arith strict-fma --tag _slow > gpurun_out/probe_strict_slow.log 2>&1; tail -2 gpurun_out/probe_strict_slow.log
python tools/pcie_probe.py > gpurun_out/pcie1.log 2>&1; tail -2 gpurun_out/pcie1.log
python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.log 2>&1; tail -2 gpurun_out/bench_ref.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_ours.log 2>&1; tail -3 gpurun_out/bench_ours.log
