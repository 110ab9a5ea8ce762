"""Example scripts stay runnable: each is started as a subprocess on the CPU path with tiny settings
(the reference's end-to-end check is "run the example and look at the printed loss")."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EX = os.path.join(ROOT, 'examples')


def _run(args, cwd=None, timeout=600, env=None):
    e = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''), OMP_NUM_THREADS='2')
    e.update(env or {})
    r = subprocess.run([sys.executable] + args, cwd=cwd or ROOT, env=e, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    return r.stdout


def test_runner_context_list_scripts():
    out = _run([os.path.join(EX, 'runner', 'run_mlp.py'), '--cpu', '--num-epochs', '2', '--samples', '2048', '--learning-rate', '0.01'])
    losses = [float(l.split('=')[1]) for l in out.splitlines() if l.startswith('Loss')]
    assert len(losses) == 2 and losses[1] < losses[0]
    out = _run([os.path.join(EX, 'runner', 'run_wdl.py'), '--cpu', '--num-epochs', '1', '--samples', '2000'])
    assert 'epoch 0 loss' in out


def test_hetuml_and_galvatron_scripts():
    out = _run([os.path.join(EX, 'hetuml', 'standalone_ml.py'), '--model', 'lr'])
    assert 'LogisticRegression valid' in out
    out = _run([os.path.join(ROOT, 'tools', 'Galvatron', 'search.py'), '--gpus', '8', '--top', '3'])
    assert 'best: heturun -w 8' in out


def test_bert_data_builder_feeds_the_trainer(tmp_path):
    corpus = tmp_path / 'corpus.txt'
    words = ['graph', 'kernel', 'tensor', 'server', 'worker', 'cache', 'token', 'expert', 'gradient', 'pipeline']
    import random
    random.seed(0)
    with open(corpus, 'w') as f:
        for _ in range(12):
            for _ in range(random.randint(3, 6)):
                f.write(' '.join(random.choice(words) for _ in range(random.randint(4, 9))) + ' .\n')
            f.write('\n')
    npz = str(tmp_path / 'train.npz')
    out = _run([os.path.join(EX, 'nlp', 'bert', 'create_pretraining_data.py'), '--input', str(corpus), '--output', npz,
                '--max-seq-length', '32', '--dupe-factor', '2'])
    assert 'wrote' in out and os.path.exists(npz)
    import numpy as np
    z = np.load(npz)
    assert z['input_ids'].shape[1] == 32 and (z['masked_lm_labels'] >= 0).any() and set(np.unique(z['next_sentence_label'])) <= {0, 1}


def test_transformer_data_pipeline(tmp_path):
    de, en = tmp_path / 'raw.de', tmp_path / 'raw.en'
    with open(de, 'w') as a, open(en, 'w') as b:
        for i in range(120):
            a.write('ich sehe haus nummer %d und baum\n' % (i % 7))
            b.write('i see house number %d and tree\n' % (i % 7))
    out_dir = str(tmp_path / 'iwslt')
    _run([os.path.join(EX, 'nlp', 'prepare_data.py'), '--src', str(de), '--tgt', str(en), '--out', out_dir, '--vocab_size', '40'])
    out = _run([os.path.join(EX, 'nlp', 'train_hetu_transformer.py'), '--data-dir', out_dir, '--maxlen', '64', '--d-model', '32',
                '--num-blocks', '1', '--batch-size', '8', '--gpu', '-1', '--epochs', '1', '--vocab-size', '40'])
    assert 'loss' in out


def _free_port_block(span=4):
    from conftest import free_port_block
    return free_port_block(span)


def test_gnn_on_sampled_subgraphs(tmp_path):
    _run(['-m', 'hetu_61a7_b200.graphmix.partition', '-d', 'Cora', '-n', '2', '-p', str(tmp_path), '--scale', '0.2'])
    out = _run([os.path.join(EX, 'gnn', 'run_dist.py'), '--data', os.path.join(str(tmp_path), 'Cora'), '--workers', '2', '--gpu', '-1',
                '--steps', '15', '--port', str(_free_port_block())])
    accs = [float(l.rsplit(' ', 1)[1]) for l in out.splitlines() if 'held-out acc' in l]
    assert accs and max(accs) > 0.6, out[-1500:]


def _torchrun(n, script, *args, port=0):
    import socket
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
    return _run(['-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(n), '--master-addr', '127.0.0.1',
                 '--master-port', str(port), script] + list(args), cwd=os.path.dirname(script),
                env={'HETU_COMM_BACKEND': 'gloo', 'OMP_NUM_THREADS': '1'})


def test_parallel_layout_examples_gloo():
    """The DP x PP, DP + MP + PP and mixed CNN layouts of examples/runner/parallel (reference scripts of the same names)
    run on a gloo CPU world; the CNN layout reproduces the single-device loss curve."""
    import numpy as np
    par = os.path.join(EX, 'runner', 'parallel')
    out = _torchrun(4, os.path.join(par, 'gpipe_multi.py'), '--steps', '2', '--hidden', '64')
    assert out.count('GPIPE_MULTI rank') == 4
    out = _torchrun(5, os.path.join(par, 'data_model_pipeline_mlp.py'), '--steps', '2', '--hidden', '64')
    assert out.count('DATA_MODEL_PIPELINE rank') == 5
    _run([os.path.join(par, 'test_model_cnn_base.py')], cwd=par, env={'HETU_COMM_BACKEND': 'gloo'})
    _torchrun(5, os.path.join(par, 'test_model_cnn_complex.py'), '--split', 'middle')
    base = np.load(os.path.join(par, 'results', 'cnn_base.npy'))
    got = np.load(os.path.join(par, 'results', 'cnn_complex_middle.npy'))
    np.testing.assert_allclose(got, base[:len(got)], rtol=1e-5, atol=1e-6)
