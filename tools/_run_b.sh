python -m pytest tests/test_ops_gpu.py -m gpu -q --maxfail=40 -p no:cacheprovider > gpurun_out/r2b_tests.log 2>&1; tail -8 gpurun_out/r2b_tests.log
python tools/conv_microbench.py resnet > gpurun_out/r2b_mb_resnet.txt 2>&1; cat gpurun_out/r2b_mb_resnet.txt
python tools/conv_microbench.py amoeba > gpurun_out/r2b_mb_amoeba.txt 2>&1; cat gpurun_out/r2b_mb_amoeba.txt
python tools/profile_step.py --model amoebanetd --batch 40 --chunks 1 --rows 60 > gpurun_out/r2b_prof_amoeba.txt 2>&1; tail -3 gpurun_out/r2b_prof_amoeba.txt
