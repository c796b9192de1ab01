#!/bin/bash
# the driver's round-end sequence on one GPU: suite, smoke, bench (ours + reference arm)
O=gpurun_out/${1:-suite}; mkdir -p $O; rm -f $O/*
timeout 900 python -m pytest tests -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log; tail -3 $O/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
O=$O python - <<'PY'
import json,sys,os; O=os.environ["O"]
for f in ('bench_n1','bench_ref'):
    for line in open('%s/%s.json' % (O, f)):
        if line.startswith('{'):
            d=json.loads(line)
            print(f, {k:d.get(k) for k in ('value','ms_per_step','ms_median_step','gpu_launches')}, 'e2e', d['e2e']['value'], 'roof', d.get('roofline',{}).get('frac'), d.get('clocks'))
            print('   api', d.get('e2e_ndarray_api',{}).get('value'), d.get('e2e_list_api',{}).get('value'), 'cpu', d.get('cpu_baseline',{}).get('value'))
PY
