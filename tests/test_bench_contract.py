"""The bench.py contract the driver depends on: the reference arm's JSON line (CPU, runs here) and -- on a GPU -- the
keys of this repo's own line."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _line(out):
    lines = [l for l in out.splitlines() if l.startswith('{')]
    assert len(lines) == 1, out[-2000:]
    return json.loads(lines[0])


def test_reference_arm_line():
    """`bench.py --impl reference`: the oracle port on the host cores, a bounded sample per pass, its real workload
    stated (rays_per_pass), e2e = value with zero copy bytes, one host process whatever --gpus says."""
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--workload', 'cfg1', '--steps', '1',
                        '--warmup', '0', '--ref-rays', '8'], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    d = _line(r.stdout)
    assert d['impl'] == 'reference' and d['metric'] == 'rays_per_sec_fwd_bwd' and d['unit'] == 'rays/s'
    assert d['higher_is_better'] is True and d['vs_baseline'] is None and d['steps'] == 1
    assert d['config']['rays_per_pass'] == 8 and d['config']['rays_per_gpu'] == 1024 and 'reference_arm' in d['config']
    assert d['cpu_baseline']['kind'] == 'port' and d['cpu_baseline']['cores'] >= 1 and d['cpu_baseline']['value'] == d['value']
    assert d['e2e'] == {'value': d['value'], 'unit': 'rays/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}
    assert d['host_processes'] == 1 and d['value'] > 0


def test_reference_arm_other_ranks_print_nothing():
    env = dict(os.environ, RANK='1', WORLD_SIZE='2')
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference', '--gpus', '2'], capture_output=True,
                       text=True, timeout=120, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ''


@pytest.mark.gpu
def test_b200_arm_line_keys():
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--workload', 'cfg1', '--steps', '3', '--warmup', '3',
                        '--no-cpu-baseline', '--no-other-configs'], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    d = _line(r.stdout)
    for k in ('metric', 'value', 'unit', 'n_gpus', 'steps', 'warmup', 'ms_per_step', 'higher_is_better', 'scaling', 'vs_baseline',
              'dtype', 'data', 'config', 'e2e', 'gpu_launches', 'clocks', 'roofline', 'stages'):
        assert k in d, k
    assert d['scaling'] == 'weak' and d['data'] == 'synthetic' and d['gpu_launches'] > 0 and d['warmup'] >= 3
    assert set(('value', 'unit', 'h2d_bytes_per_step', 'd2h_bytes_per_step')) <= set(d['e2e']) and d['e2e']['h2d_bytes_per_step'] > 0
    assert set(('bound', 'achieved', 'peak', 'unit', 'frac', 'traffic')) <= set(d['roofline'])
    assert 'workload' in d['config'] and 'model' not in d['config']
    assert set(('sm_mhz', 'sm_max_mhz', 'reasons')) <= set(d['clocks'])
