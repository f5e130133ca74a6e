# 2 GPUs: multi-GPU correctness, pipelined e2e, layouts, PCIe ceiling
set -x
nvidia-smi topo -m > gpurun_out/topo_n2.txt 2>&1
timeout 900 python -m pytest tests/test_dist_gpu.py tests/test_gpu_layouts.py -m gpu -q --timeout=300 > gpurun_out/gputest_multi_n2.log 2>&1; echo "rc=$?" >> gpurun_out/gputest_multi_n2.log; tail -8 gpurun_out/gputest_multi_n2.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29601 tools/pcie_probe.py > gpurun_out/pcie2.log 2>&1; tail -1 gpurun_out/pcie2.log
timeout 300 $TR --master-port 29602 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_n2.log 2>&1; tail -1 gpurun_out/bench_n2.log
timeout 300 $TR --master-port 29603 bench.py --gpus 2 --steps 1000 --warmup 24 --arith fp32 --no-extra --px 2 --py 1 > gpurun_out/bench_n2_fp32_2x1.log 2>&1; tail -1 gpurun_out/bench_n2_fp32_2x1.log | cut -c1-400
timeout 300 $TR --master-port 29604 bench.py --gpus 2 --steps 1000 --warmup 24 --arith fp32 --no-extra --no-e2e --px 1 --py 2 > gpurun_out/bench_n2_fp32_1x2.log 2>&1; tail -1 gpurun_out/bench_n2_fp32_1x2.log | cut -c1-400
timeout 300 $TR --master-port 29605 bench.py --gpus 2 --steps 1000 --warmup 24 --arith fp32 --no-extra --no-e2e --impl nccl-halo --k 8 > gpurun_out/bench_n2_nccl_k8.log 2>&1; tail -1 gpurun_out/bench_n2_nccl_k8.log | cut -c1-400
