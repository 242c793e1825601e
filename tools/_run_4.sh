cd benchmarks
for e in baseline dep-x-x dep-str-x dep-str-ptl product; do timeout 200 python timeline.py $e --steps 5 --warmup 3 2>/dev/null | grep '^{' | tail -1 >> ../gpurun_out/r2_timeline.jsonl; done
cut -c1-260 ../gpurun_out/r2_timeline.jsonl
cd ..
(time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 --steps 5 --warmup 3) > gpurun_out/r2_ours4.json 2> gpurun_out/r2_ours4.err; tail -c 2500 gpurun_out/r2_ours4.json | cut -c1-2500; tail -3 gpurun_out/r2_ours4.err
(time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 4 --steps 3 --warmup 3 --impl reference) > gpurun_out/r2_ref4.json 2> gpurun_out/r2_ref4.err; tail -c 1200 gpurun_out/r2_ref4.json; tail -3 gpurun_out/r2_ref4.err
python benchmarks/speed.py unet pipeline-4 --graphs --steps 5 2>/dev/null | tail -1 | cut -c1-300
