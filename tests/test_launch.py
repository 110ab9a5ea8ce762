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


@pytest.mark.parametrize('mode,policy', [('PS', 'none'), ('Hybrid', 'lfuopt')])
def test_checkpoint_round_trip_in_ps_modes(mode, policy):
    """save -> drift -> load -> the next step reproduces the step taken right after the save (servers' copies of dense
    and embedding parameters restored, replicas refreshed, HET cache dropped)."""
    with tempfile.TemporaryDirectory() as d:
        env = dict(os.environ, HETU_CONFIG_PATH=os.path.join(d, 'hetu_config.yml'), OMP_NUM_THREADS='2')
        cmd = [os.path.join(ROOT, 'bin', 'heturun'), '-s', '1', '-w', '2', sys.executable,
               os.path.join(ROOT, 'tests', 'ps_ckpt_worker.py'), mode, policy, d]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0 and r.stdout.count('OK') == 2, r.stdout[-2000:] + r.stderr[-3000:]


def test_heturun_detects_a_dead_worker_and_restarts(tmp_path):
    """A worker that exits with an error takes the group down instead of leaving the others hanging in a
    collective; with --max-restarts the group is relaunched and resumes from its checkpoint."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    worker = os.path.join(root, 'tests', 'restart_worker.py')
    env = dict(os.environ, HETU_CONFIG_PATH=str(tmp_path / 'cfg.yml'), OMP_NUM_THREADS='2')
    d1 = tmp_path / 'a'
    d1.mkdir()
    r = subprocess.run([sys.executable, '-m', 'hetu_61a7_b200.runner', '-w', '2', sys.executable, worker, str(d1)], cwd=root, env=env,
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 17, r.stdout[-2000:] + r.stderr[-2000:]          # no restarts: the failure is reported, nothing hangs
    d2 = tmp_path / 'b'
    d2.mkdir()
    r = subprocess.run([sys.executable, '-m', 'hetu_61a7_b200.runner', '-w', '2', '--max-restarts', '1', sys.executable, worker, str(d2)],
                       cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert 'restarting all workers (1 of 1)' in r.stdout
    assert r.stdout.count('incarnation 1 finished from step 3') == 2
