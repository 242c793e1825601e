timeout 600 python -m pytest tests/test_graph_pipeline_gpu.py -m gpu -q --maxfail=10 -p no:cacheprovider > gpurun_out/r2c_tests.log 2>&1; tail -40 gpurun_out/r2c_tests.log
