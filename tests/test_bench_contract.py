"""bench.py's reference arm runs without a GPU: check that it prints exactly one
JSON line on stdout with the keys the driver reads, and nothing else there."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference',
                          '--steps', '2', '--warmup', '1', '--cpu-size', '256'],
                         capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert j['impl'] == 'reference' and j['unit'] == 'cell-updates/s' and j['higher_is_better'] is True
    assert j['metric'] == 'grid cell-updates/sec (fp64 routing step)'
    assert j['steps'] == 2 and j['warmup'] == 1 and j['n_gpus'] == 1 and j['vs_baseline'] is None
    assert j['value'] > 0 and abs(j['e2e']['value'] - j['value']) < 1e-6 * j['value']
    assert j['e2e']['h2d_bytes_per_step'] == 0 and j['e2e']['d2h_bytes_per_step'] == 0
    assert j['cpu_baseline']['kind'] == 'port' and j['cpu_baseline']['cores'] == 1
    assert j['config']['workload'].startswith('synthetic 16384x16384') and 'sample' in j['config']


def test_other_ranks_of_the_reference_arm_do_nothing():
    env = dict(os.environ, RANK='1', WORLD_SIZE='2', LOCAL_RANK='1')
    out = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--impl', 'reference',
                          '--gpus', '2', '--steps', '2', '--warmup', '1', '--cpu-size', '256'],
                         capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ''
