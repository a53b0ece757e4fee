"""The driver-facing contract of bench.py and __graft_entry__.py that can be checked without a GPU."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), *args], capture_output=True, text=True, timeout=300,
                       env={**os.environ, 'CUDA_VISIBLE_DEVICES': ''})
    lines = [l for l in r.stdout.splitlines() if l.startswith('{')]
    return r.returncode, lines


def test_reference_arm_prints_one_unavailable_line_and_exits_zero():
    rc, lines = _run('--impl', 'reference', '--gpus', '1', '--steps', '5', '--warmup', '3')
    assert rc == 0 and len(lines) == 1
    d = json.loads(lines[0])
    assert d['impl'] == 'reference' and 'unavailable' in d and 'pyspark' in d['unavailable']


def test_bench_without_a_gpu_says_so_in_one_json_line():
    rc, lines = _run('--steps', '5', '--warmup', '3')
    assert rc == 0 and len(lines) == 1
    assert json.loads(lines[0]).get('unavailable') == 'no CUDA device visible'


def test_graft_entry_exposes_build_and_smoke():
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    assert callable(g.build) and callable(g.smoke)
    g.build()                                            # nvcc cross-compiles for sm_100a without a GPU (cached objects)
    assert os.path.exists(os.path.join(ROOT, 'distributeddeeplearningspark_b200', '_lib', 'libddls_b200.so'))
