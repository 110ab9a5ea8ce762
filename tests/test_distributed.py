"""Multi-process tests: gloo world of 2 on CPU (host-side logic) and, on a multi-GPU box, the
in-kernel NVLink collectives / fused DP update / sharded embedding / fused MoE exchange against
the NCCL baseline (pattern: reference examples/runner/parallel/validate_results.py)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def _launch(mode, nproc, timeout=600):
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(nproc),
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.join(ROOT, 'tests', 'dist_worker.py'), mode]
    env = dict(os.environ, OMP_NUM_THREADS='2')
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
    assert r.returncode == 0 and 'ALL DIST CHECKS PASSED' in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_gloo_world2_host_logic():
    _launch('cpu', 2)


@pytest.mark.gpu
@pytest.mark.timeout(900)
def test_multi_gpu_symm_paths():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip('needs >= 2 GPUs (gpurun --gpus 2)')
    _launch('gpu', min(n, 4), timeout=800)


def _launch_pp(mode, case, timeout=600, nproc=2):
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(nproc),
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.join(ROOT, 'tests', 'pp_worker.py'), mode, case]
    import tempfile
    with tempfile.TemporaryDirectory() as ck:           # gpipe / left also do a per-rank checkpoint round trip
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout,
                           env=dict(os.environ, OMP_NUM_THREADS='2', HETU_TEST_CKPT_DIR=ck))
    want = 4 if case in ('gpipe', 'left') else nproc
    assert r.returncode == 0 and r.stdout.count('OK') == want, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize('case', ['gpipe', 'pipedream', 'left', 'right', 'middle', 'sp'])
def test_pipeline_and_model_parallel_equivalence_gloo(case):
    _launch_pp('cpu', case)


def test_pipeline_with_unequal_data_parallel_degrees_gloo():
    """Stage 0 on one rank, stage 1 on two replicas (reference pipeline_subexecutor.py:83-106)."""
    _launch_pp('cpu', 'unequal', nproc=3)


@pytest.mark.gpu
@pytest.mark.timeout(900)
@pytest.mark.parametrize('case', ['gpipe', 'pipedream', 'left', 'right', 'middle', 'sp'])
def test_pipeline_and_model_parallel_equivalence_nccl(case):
    if torch.cuda.device_count() < 2:
        pytest.skip('needs >= 2 GPUs (gpurun --gpus 2)')
    _launch_pp('gpu', case)


def _launch_hybrid(mode, pp, tp, dp, pipeline=None, timeout=900):
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(pp * tp * dp),
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.join(ROOT, 'tests', 'bert_hybrid_worker.py'), mode, str(pp), str(tp), str(dp)]
    if pipeline:
        cmd.append(pipeline)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.timeout(900)
@pytest.mark.parametrize('pp,tp,dp,pipeline', [(1, 2, 1, None), (2, 2, 1, 'gpipe'), (2, 2, 2, None)])
def test_bert_hybrid_parallel_gloo(pp, tp, dp, pipeline):
    """BERT with pipeline stages x tensor-parallel FFN x data-parallel replicas == single process
    (BASELINE config 5 layout at toy size)."""
    _launch_hybrid('cpu', pp, tp, dp, pipeline)


@pytest.mark.gpu
@pytest.mark.timeout(900)
def test_bert_hybrid_parallel_nccl():
    n = torch.cuda.device_count()
    if n < 4:
        pytest.skip('needs >= 4 GPUs')
    if n >= 8:
        _launch_hybrid('gpu', 2, 2, 2, 'gpipe')
    else:
        _launch_hybrid('gpu', 2, 2, 1, 'gpipe')
