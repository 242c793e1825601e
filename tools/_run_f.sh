cd benchmarks
for e in baseline dep-x-x dep-str-x dep-str-ptl; do timeout 200 python timeline.py $e --steps 5 --warmup 3 2>/dev/null | grep '^{' | tail -1 >> ../gpurun_out/r2f_timeline.jsonl; done
cat ../gpurun_out/r2f_timeline.jsonl | cut -c1-300
for e in baseline pipeline-1 pipeline-2 pipeline-4; do timeout 240 python memory.py amoebanetd $e 2>/dev/null | grep '^{' | tail -1 >> ../gpurun_out/r2f_memory_amoebanetd.jsonl; done
cat ../gpurun_out/r2f_memory_amoebanetd.jsonl | cut -c1-300
for e in baseline pipeline-1 pipeline-2 pipeline-4; do (time timeout 400 python memory.py unet $e 2>/dev/null | grep '^{' | tail -1 >> ../gpurun_out/r2f_memory_unet.jsonl) 2>&1 | grep real; done
cat ../gpurun_out/r2f_memory_unet.jsonl | cut -c1-300
