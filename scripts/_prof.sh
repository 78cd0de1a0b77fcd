GBL_USE_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -s 330 -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
GBL_USE_GRAPH=0 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:hist_float -s 54 -c 36 --csv --log-file gpurun_out/hist_traffic.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_traffic.log 2>&1
tail -2 gpurun_out/ncu_list.log | cut -c1-300
