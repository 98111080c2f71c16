#!/bin/bash
# One short GPU session: selected parity tests, then the default bench line with and without one diagnostic switch.
# usage: gpu_call_ab.sh TAG ENVVAR "pytest -k expression"
tag=${1:-cXX}; var=$2; kexpr=$3
timeout 240 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -s -k "$kexpr" > gpurun_out/${tag}_tests.log 2>&1; rc=$?
echo "rc=$rc" >> gpurun_out/${tag}_tests.log
grep -E "passed|failed|rc=|Error|assert" gpurun_out/${tag}_tests.log | tail -12
[ $rc -eq 0 ] || exit 1
for v in 0 1 0; do
  env $var=$v timeout 200 python bench.py --no-extra --no-cpu --steps 6 --warmup 3 --layers \
      > gpurun_out/${tag}_bench_$var$v.json 2> gpurun_out/${tag}_bench_$var$v.err || { tail -5 gpurun_out/${tag}_bench_$var$v.err; exit 1; }
  python -c "import json;d=json.load(open('gpurun_out/${tag}_bench_$var$v.json'));print('$var=$v', round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'], round(d['roofline']['frac'],4), d.get('cm_sha256','')[:12], d['gpu_launches'])"
done
