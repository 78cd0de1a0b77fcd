(time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r2_bench_8gpu.log 2>gpurun_out/r2_bench_8gpu.err) 2>&1 | tail -3
tail -c 600 gpurun_out/r2_bench_8gpu.err
(time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r2_bench_4gpu.log 2>gpurun_out/r2_bench_4gpu.err) 2>&1 | tail -3
