python tools/_dbg2.py cuda:0,cuda:1 2>&1 | grep -v Warning | grep "True" | head
for i in 1 2; do timeout 300 python -m pytest tests/test_graph_pipeline_gpu.py -m gpu -q -p no:cacheprovider 2>&1 | tail -2; done
(time timeout 900 python -m pytest tests -m gpu -q --maxfail=10 -p no:cacheprovider) > gpurun_out/r2d_tests.log 2>&1; tail -4 gpurun_out/r2d_tests.log
python benchmarks/speed.py amoebanetd n2m4 --graphs --steps 5 2>&1 | tail -1
python benchmarks/speed.py amoebanetd n2m4 --steps 5 2>&1 | tail -1
python benchmarks/speed.py amoebanetd n2m4 --impl reference --steps 5 2>&1 | tail -1
