#!/bin/bash
# One GPU session for the CLIP ViT: tests, timings, ncu launch list, ClipUnet bench per CLIP backend.  Outputs: gpurun_out/$1_*
tag=${1:-cXX}
timeout 600 python -m pytest tests/test_gpu_clip_vit.py -x -q -m gpu -s > gpurun_out/${tag}_clip_vit.log 2>&1; rc=$?
echo "rc=$rc" >> gpurun_out/${tag}_clip_vit.log
grep -v Warning gpurun_out/${tag}_clip_vit.log | tail -14
[ $rc -eq 0 ] || exit 1
timeout 300 python tools/time_clip_vit.py 122 > gpurun_out/${tag}_time_vit_122.txt 2>&1; tail -9 gpurun_out/${tag}_time_vit_122.txt
timeout 300 python tools/time_clip_vit.py 244 > gpurun_out/${tag}_time_vit_244.txt 2>&1; head -3 gpurun_out/${tag}_time_vit_244.txt
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_vit_launches.csv \
    python tools/time_clip_vit.py 122 ours > gpurun_out/${tag}_ncu_vit.log 2>&1
for be in b200 transformers-bf16 transformers; do
  UB_CLIP_BACKEND=$be timeout 600 python bench.py --workload clipunet256 --no-extra --steps 3 --warmup 3 \
      > gpurun_out/${tag}_bench_clipunet_$be.json 2> gpurun_out/${tag}_bench_clipunet_$be.err
  python -c "import json;d=json.load(open('gpurun_out/${tag}_bench_clipunet_$be.json'));print('$be', d['value'], d['e2e']['value'], d['clocks']['sm_mhz'])"
done
