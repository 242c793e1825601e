(time python bench.py --steps 10 --warmup 3) > gpurun_out/r2c_ours.json 2> gpurun_out/r2c_ours.err; tail -c 1800 gpurun_out/r2c_ours.json; tail -5 gpurun_out/r2c_ours.err
(time python bench.py --steps 10 --warmup 3 --model resnet101 --no-secondary) > gpurun_out/r2c_resnet.json 2> gpurun_out/r2c_resnet.err; tail -c 600 gpurun_out/r2c_resnet.json; tail -5 gpurun_out/r2c_resnet.err
(time python bench.py --steps 5 --warmup 3 --model resnet101 --impl reference --no-secondary) > gpurun_out/r2c_resnet_ref.json 2> gpurun_out/r2c_resnet_ref.err; tail -c 400 gpurun_out/r2c_resnet_ref.json
python tools/profile_step.py --graphs --rows 50 > gpurun_out/r2c_prof_unet.txt 2>&1; tail -3 gpurun_out/r2c_prof_unet.txt
