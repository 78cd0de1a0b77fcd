# Usage: bash scripts/run_multi_gpu.sh N   (N = GPUs of one node): N-GPU trees == 1-GPU trees, then the bench at N GPUs on both workloads
N=$1
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scripts/check_multi_gpu.py > gpurun_out/check_mg$N.log 2>&1
echo "check rc=$?"; tail -3 gpurun_out/check_mg$N.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${N}gpu_peer.log 2>&1
echo "bench rc=$?"; tail -1 gpurun_out/bench_${N}gpu_peer.log | cut -c1-250
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --workload reg100m --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_reg_${N}gpu.log 2>&1
echo "reg rc=$?"; tail -1 gpurun_out/bench_reg_${N}gpu.log | cut -c1-250
