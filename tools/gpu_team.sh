set -x
timeout 600 python -m pytest tests/test_gpu_team.py -m gpu -x -q --timeout=120 > gpurun_out/gputest_team.log 2>&1; echo "rc=$?" >> gpurun_out/gputest_team.log; tail -15 gpurun_out/gputest_team.log
timeout 300 python tools/team_probe.py --n 32768 --launches 50 2>&1 | grep -v Warning
