"""heturun end-to-end: scheduler + server + 2 workers, PS and Hybrid (+HET cache) training on CPU
(pattern: reference examples/cnn/scripts/hetu_2gpu_ps.sh / examples/ctr/tests/hybrid_wdl_criteo.sh)."""
import os
import subprocess
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize('mode,policy,extra', [('PS', 'none', ''), ('Hybrid', 'lfuopt', ''), ('PS', 'none', 'prefetch')])
def test_heturun_ps_modes(mode, policy, extra):
    with tempfile.TemporaryDirectory() as d:
        env = dict(os.environ, HETU_CONFIG_PATH=os.path.join(d, 'hetu_config.yml'), OMP_NUM_THREADS='2')
        cmd = [os.path.join(ROOT, 'bin', 'heturun'), '-s', '1', '-w', '2', sys.executable,
               os.path.join(ROOT, 'tests', 'ps_train_worker.py'), mode, policy, d] + ([extra] if extra else [])
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0 and r.stdout.count('OK') == 2, r.stdout[-2000:] + r.stderr[-3000:]
        files = os.listdir(d)
        assert 'ckpt_%s.bin' % mode in files and any(f.endswith('.dat') for f in files), files
