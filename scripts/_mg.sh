export GBL_DEBUG_PEER=1
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/check_multi_gpu.py > gpurun_out/check_mg.log 2>&1
echo "check rc=$?"; grep -c "exchange on" gpurun_out/check_mg.log; grep -c "OFF" gpurun_out/check_mg.log; tail -3 gpurun_out/check_mg.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_2gpu_peer.log 2>&1
echo "bench rc=$?"; grep "gbl\]" gpurun_out/bench_2gpu_peer.log; tail -1 gpurun_out/bench_2gpu_peer.log | cut -c1-250
