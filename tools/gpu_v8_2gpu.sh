# gpurun payload (2 GPUs): two-GPU correctness cases + the driver's bench command on 2 GPUs
set -x
timeout 200 python -m pytest tests/test_dist_gpu.py -m gpu -q --timeout=180 -x -k "two" 2>&1 | tail -2
bash tools/gpu_bench_n.sh 2
