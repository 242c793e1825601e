set -x
nvidia-smi --query-gpu=name,memory.total --format=csv,noheader
(time python -m pytest tests -m gpu -x -q) > gpurun_out/r2a_tests.log 2>&1; tail -5 gpurun_out/r2a_tests.log
(time python bench.py --impl reference --steps 5 --warmup 3) > gpurun_out/r2a_ref.json 2> gpurun_out/r2a_ref.err; tail -c 600 gpurun_out/r2a_ref.json
(time python bench.py --steps 10 --warmup 3) > gpurun_out/r2a_ours.json 2> gpurun_out/r2a_ours.err; tail -c 1500 gpurun_out/r2a_ours.json; tail -5 gpurun_out/r2a_ours.err
build/selftest unetb16 > gpurun_out/r2a_selftest_mb16.txt 2>&1; tail -3 gpurun_out/r2a_selftest_mb16.txt
build/selftest unet > gpurun_out/r2a_selftest_mb40.txt 2>&1; tail -3 gpurun_out/r2a_selftest_mb40.txt
