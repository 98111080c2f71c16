#!/bin/bash
# Final check of a build: full GPU suite, smoke(), the default bench line, then (optional, "ncu") the --set full capture of one
# micro-batch.  Outputs: gpurun_out/$1_*
tag=${1:-cXX}
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_pytest_gpu.log 2>&1; rc=$?
echo "rc=$rc" >> gpurun_out/${tag}_pytest_gpu.log
grep -E "passed|failed|rc=" gpurun_out/${tag}_pytest_gpu.log | tail -3
[ $rc -eq 0 ] || { tail -30 gpurun_out/${tag}_pytest_gpu.log; exit 1; }
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1 || { tail -5 gpurun_out/${tag}_smoke.log; exit 1; }
tail -1 gpurun_out/${tag}_smoke.log
timeout 400 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || { tail -5 gpurun_out/${tag}_bench.err; exit 1; }
python -c "import json;d=json.load(open('gpurun_out/${tag}_bench.json'));print(round(d['value']), round(d['e2e']['value']), d['clocks']['sm_mhz'], round(d['roofline']['frac'],4), d.get('cm_sha256','')[:12], d['gpu_launches'], d.get('extra'))"
if [ "$2" = "ncu" ]; then
  timeout 400 ncu --set full --clock-control none -k regex:"conv_tc|conv_rs2|image_pack|resample" -s 19 -c 19 -o /tmp/final_prof \
      python tools/run_forward_once.py 148 > gpurun_out/${tag}_ncu.log 2>&1; tail -1 gpurun_out/${tag}_ncu.log
  ncu -i /tmp/final_prof.ncu-rep --page raw --csv > gpurun_out/${tag}_final_raw.csv 2>/dev/null; ls -la gpurun_out/${tag}_final_raw.csv
fi
