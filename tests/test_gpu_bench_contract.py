"""The bench harness end to end on the GPU: `bench.py --steps 2 --warmup 1 --no-cpu-baseline` must print one
JSON line with the keys the measurement contract names, measured live (not read from a committed file)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bench_prints_a_live_line_with_the_contract_keys():
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--steps', '2', '--warmup', '1', '--no-cpu-baseline'],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith('{')]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ('metric', 'value', 'unit', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better', 'scaling',
              'vs_baseline', 'dtype', 'data', 'config', 'clocks', 'e2e', 'gpu_launches', 'roofline'):
        assert k in d, k
    assert d['n_gpus'] == 1 and d['steps'] == 2 and d['value'] > 0 and d['vs_baseline'] is None
    assert d['gpu_launches'] >= 2                       # counted by the library, not assumed
    assert d['e2e']['value'] > 0 and d['e2e']['h2d_bytes_per_step'] == 65536 * 784
    roof = d['roofline']
    assert roof['bound'] == 'tensor' and 0 < roof['frac'] < 1.5 and roof['frac_burst'] <= roof['frac_sustained']
    assert abs(roof['achieved'] / roof['peak'] - roof['frac']) < 1e-9
    assert 'workload' in d['config'] and 'model' not in d['config']
