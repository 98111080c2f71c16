#!/bin/bash
# One short GPU session for the ConvTranspose+resize epilogue: its tests, then the default bench line with and without it.
tag=${1:-cXX}
timeout 420 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -s \
    -k "one_epilogue or FUSE_RESIZE or large_unet_forward or forward_logits_and_argmax or segmentation_model or encoder_and_skip" > gpurun_out/${tag}_resize_tests.log 2>&1; rc=$?
echo "rc=$rc" >> gpurun_out/${tag}_resize_tests.log
grep -E "ConvT\+resize|passed|failed|rc=|Error|assert" gpurun_out/${tag}_resize_tests.log | tail -16
[ $rc -eq 0 ] || exit 1
for v in 0 1 0; do
  UB_NO_FUSE_RESIZE=$v timeout 200 python bench.py --no-extra --no-cpu --steps 6 --warmup 3 --layers \
      > gpurun_out/${tag}_bench_nofuse$v.json 2> gpurun_out/${tag}_bench_nofuse$v.err || { tail -5 gpurun_out/${tag}_bench_nofuse$v.err; exit 1; }
  python -c "import json;d=json.load(open('gpurun_out/${tag}_bench_nofuse$v.json'));print('UB_NO_FUSE_RESIZE=$v', round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'], d['roofline']['frac'], d.get('cm_sha256','')[:12], d['gpu_launches'])"
done
grep -E "dec1|resample" gpurun_out/${tag}_bench_nofuse0.err | head; grep -E "dec1|resample" gpurun_out/${tag}_bench_nofuse1.err | head
