#!/bin/bash
# round-end evidence on one GPU: suite, smoke, bench lines (headline, reference arm, small, c2, c4, c5), ncu of the small kernels
O=gpurun_out/${1:-final}; mkdir -p $O; rm -f $O/*
timeout 900 python -m pytest tests -q -m gpu -x > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $O/smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
timeout 300 python bench.py --workload small --steps 5 --warmup 3 > $O/bench_small.json 2> $O/bench_small.err; echo "small rc=$?"
timeout 300 python bench.py --workload c2 --steps 200 --warmup 10 > $O/bench_c2.json 2> $O/bench_c2.err; echo "c2 rc=$?"
timeout 300 python bench.py --workload c4 --steps 50 --warmup 5 > $O/bench_c4.json 2> $O/bench_c4.err; echo "c4 rc=$?"
timeout 600 python bench.py --workload c5 --steps 200 > $O/bench_c5.json 2> $O/bench_c5.err; echo "c5 rc=$?"
SM="python bench.py --workload small --steps 5 --warmup 1"
ncu --set full --clock-control none -k regex:'decode_unit|resize_u8_staged|ensemble_mean_vec4' -s 4 -c 12 -o $O/small $SM > $O/ncu_small.log 2>&1
echo "small capture rc=$?"
python - <<PY
import json
for f in ('bench_n1','bench_ref','bench_small','bench_c2','bench_c4','bench_c5'):
    try:
        d=json.loads([l for l in open('$O/%s.json' % f) if l.startswith('{')][-1])
        print(f, 'value', d.get('value'), 'ms/step', d.get('ms_per_step'), 'roof', (d.get('roofline') or {}).get('frac'), 'e2e', (d.get('e2e') or {}).get('value'))
    except Exception as e:
        print(f, 'ERR', e)
PY
