"""Data-parallel throughput of the secondary BASELINE configs (ResNet-18 / CIFAR10 shape, BERT-large seq 128):
device-timed, max over ranks, one JSON line from rank 0.
usage: [torchrun --nproc-per-node N] dp_bench.py {resnet18|bert_large|bert_base|mlp_cpu} [steps]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hetu as ht

name = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
rs = np.random.RandomState(int(os.environ.get('RANK', 0)))
if name == 'mlp_cpu':                                   # BASELINE config 1: CPU executor, single process
    from hetu_61a7_b200.models import cnn
    x, y_ = ht.Variable('x'), ht.Variable('y_')
    loss, _ = cnn.mlp(x, y_)
    train = ht.optim.SGDOptimizer(0.01).minimize(loss)
    ex = ht.Executor([loss, train], ctx=ht.cpu(0))
    X = rs.randn(128, 3072).astype(np.float32)
    Y = np.eye(10, dtype=np.float32)[rs.randint(0, 10, 128)]
    for _ in range(3):
        ex.run(feed_dict={x: X, y_: Y})
    t0 = time.perf_counter()
    for _ in range(steps):
        ex.run(feed_dict={x: X, y_: Y})[0].asnumpy()
    dt = (time.perf_counter() - t0) / steps
    print(json.dumps({'config': 'mlp 3072-256-256-10, batch 128, CPU executor (OpenMP op library)', 'ms_per_step': dt * 1e3,
                      'samples_per_s': 128 / dt}))
    sys.exit(0)

from hetu_61a7_b200.parallel.comm import device_max_ms
comm = ht.wrapped_mpi_nccl_init()
world = comm.nrank
ctx = ht.gpu(comm.local_rank)
mode = 'AllReduce' if world > 1 else None
if name == 'resnet18':
    from hetu_61a7_b200.models import cnn
    x, y_ = ht.Variable('x'), ht.Variable('y_')
    loss, _ = cnn.resnet18(x, y_)
    train = ht.optim.SGDOptimizer(0.01 / world).minimize(loss)
    ex = ht.Executor([loss, train], ctx=ctx, comm_mode=mode)
    fd = {x: ht.array(rs.randn(128, 3072).astype(np.float32), ctx), y_: ht.array(np.eye(10, dtype=np.float32)[rs.randint(0, 10, 128)], ctx)}
    units, unit = 128, 'samples/s'
else:
    from hetu_61a7_b200.models.nlp import BertConfig, BertForPreTraining
    cfg = (BertConfig.large if name == 'bert_large' else BertConfig.base)(batch_size=32, seq_len=128, attention_probs_dropout_prob=0.0)
    ids, tt, am, mlm, nsp = [ht.Variable(n) for n in ('ids', 'tt', 'am', 'mlm', 'nsp')]
    _, _, l1, l2 = BertForPreTraining(cfg)(ids, tt, am, mlm, nsp)
    loss = l1 + l2
    train = ht.optim.AdamOptimizer(1e-5).minimize(loss)
    ex = ht.Executor([loss, train], ctx=ctx, comm_mode=mode)
    lab = rs.randint(0, cfg.vocab_size, (32, 128)).astype(np.float32)
    lab[rs.rand(32, 128) > 0.15] = -1
    fd = {ids: ht.array(rs.randint(0, cfg.vocab_size, (32, 128)).astype(np.float32), ctx), tt: ht.array(np.zeros((32, 128), np.float32), ctx),
          am: ht.array(np.ones((32, 128), np.float32), ctx), mlm: ht.array(lab, ctx), nsp: ht.array(rs.randint(0, 2, 32).astype(np.float32), ctx)}
    units, unit = 32 * 128, 'tokens/s'
for _ in range(4):
    ex.run(feed_dict=fd)
comm.barrier(); torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(steps):
    out = ex.run(feed_dict=fd)
e.record(); torch.cuda.synchronize()
ms = device_max_ms(s.elapsed_time(e) / steps)
l = float(out[0].asnumpy())
if comm.rank == 0:
    print(json.dumps({'config': name, 'n_gpus': world, 'ms_per_step': ms, 'value': units * world / (ms / 1e3), 'unit': unit,
                      'loss': l, 'cuda_graph': ex.subexecutor['default']._graph is not None if 'default' in ex.subexecutor else None}))
