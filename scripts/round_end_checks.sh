# Round-end checks on a GPU box: the -m gpu suite, smoke(), both bench arms (outputs under gpurun_out/).
# Make this the LAST action of a round, after the last code commit.
(time timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/gpu_tests_final.log 2>&1; tail -3 gpurun_out/gpu_tests_final.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 300 python bench.py --impl reference > gpurun_out/bench_ref.log 2>&1; tail -1 gpurun_out/bench_ref.log | cut -c1-300
(time timeout 600 python bench.py) > gpurun_out/bench_final.log 2>&1; grep '^{' gpurun_out/bench_final.log | cut -c1-200; tail -4 gpurun_out/bench_final.log | grep real
