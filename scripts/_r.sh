python scripts/run_shapes.py higgs 8 "bagging_fraction=0.9 bagging_freq=1" > gpurun_out/higgs_bag.log 2>&1
python scripts/run_shapes.py higgs 6 "metric=auc,binary_logloss,binary_error" > gpurun_out/higgs_eval.log 2>&1
python scripts/run_shapes.py mslr 6 > gpurun_out/mslr.log 2>&1
grep -E "iter|profile|OK" gpurun_out/higgs_bag.log gpurun_out/higgs_eval.log gpurun_out/mslr.log
