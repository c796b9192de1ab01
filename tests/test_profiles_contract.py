"""The committed bench lines carry every key of the driver contract (bench.py's docstring), so a change to
bench.py that drops one is caught on CPU."""
import json
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE = ['metric', 'value', 'unit', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better', 'scaling',
        'vs_baseline', 'dtype', 'data', 'config', 'clocks', 'e2e', 'gpu_launches', 'roofline']


def _line(name):
    path = os.path.join(ROOT, 'profiles', name)
    if not os.path.exists(path):
        pytest.skip(name + ' not committed')
    return json.loads(open(path).read().strip().splitlines()[-1])


@pytest.mark.parametrize('name,n', [('r01_bench_c3.json', 1), ('r01_bench_c3_n2.json', 2), ('r01_bench_c3_n8.json', 8),
                                    ('r01_bench_c2.json', 1), ('r01_bench_c4.json', 1)])
def test_bench_line_has_the_contract_keys(name, n):
    d = _line(name)
    for k in BASE:
        assert k in d, k
    assert d['n_gpus'] == n and d['value'] > 0 and d['higher_is_better'] is True and d['vs_baseline'] is None
    assert 'workload' in d['config'] and 'model' not in d['config']
    assert {'value', 'unit', 'h2d_bytes_per_step', 'd2h_bytes_per_step'} <= set(d['e2e'])
    assert d['e2e']['h2d_bytes_per_step'] > 0 and d['e2e']['value'] != d['value']
    assert {'bound', 'achieved', 'peak', 'unit', 'frac', 'traffic'} <= set(d['roofline'])
    assert abs(d['roofline']['frac'] - d['roofline']['achieved'] / d['roofline']['peak']) < 1e-9
    assert {'sm_mhz', 'sm_max_mhz', 'reasons'} <= set(d['clocks'])
    assert d['gpu_launches'] >= d['steps']
    assert not ({'hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown'} & set(d['clocks']['reasons']))


def test_headline_line_carries_the_cpu_baseline_and_the_reference_arm_its_marker():
    d = _line('r01_bench_c3.json')
    assert {'value', 'unit', 'cores', 'kind', 'sample'} <= set(d['cpu_baseline']) and d['cpu_baseline']['kind'] == 'port'
    r = _line('r01_bench_reference_arm.json')
    assert r['impl'] == 'reference' and r['metric'] == d['metric'] and r['unit'] == d['unit']
    assert r['e2e']['h2d_bytes_per_step'] == 0 and r['e2e']['d2h_bytes_per_step'] == 0
    assert r['cpu_baseline']['value'] == r['value']
