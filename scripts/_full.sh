GBL_USE_GRAPH=0 timeout 500 ncu --set full --clock-control none --import-source on -k regex:hist_float -s 54 -c 3 -o gpurun_out/hist_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "rc=$?"; ls -la gpurun_out/hist_full.ncu-rep; tail -2 gpurun_out/ncu_full.log | cut -c1-200
