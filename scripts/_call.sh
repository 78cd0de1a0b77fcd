GBL_TILE_SCAN=1 timeout 900 python -m pytest tests/test_learner_parity.py tests/test_wave_schedule.py tests/test_booster_parity.py tests/test_golden_gpu.py -m gpu -x -q 2>&1 | tail -3
for ts in 1; do echo "== epsilon GBL_TILE_SCAN=$ts"; GBL_TILE_SCAN=$ts python scripts/run_shapes.py epsilon 5 2>&1 | grep -E "iter 4|profile"; done
