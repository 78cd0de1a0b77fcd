timeout 900 python -m pytest tests/test_full_size.py tests/test_step_time.py tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -30
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
(time python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench2.log 2>gpurun_out/r2_bench2.err) 2>&1 | tail -4
tail -c 600 gpurun_out/r2_bench2.err
