"""Run a few training steps of a zoo model (target of `ncu --metrics gpu__time_duration.sum` breakdowns).
usage: model_step.py {resnet18|bert_base|bert_large} [steps]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import hetu as ht
name = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
os.environ.setdefault('HETU_USE_CUDA_GRAPH', '1')
rs = np.random.RandomState(0)
ctx = ht.gpu(0)
if name == 'resnet18':
    from hetu_61a7_b200.models import cnn
    x, y_ = ht.Variable('x'), ht.Variable('y_')
    loss, y = cnn.resnet18(x, y_)
    train = ht.optim.SGDOptimizer(0.01).minimize(loss)
    ex = ht.Executor([loss, train], ctx=ctx)
    fd = {x: ht.array(rs.randn(128, 3072).astype(np.float32), ctx), y_: ht.array(np.eye(10, dtype=np.float32)[rs.randint(0, 10, 128)], ctx)}
    units = 128
else:
    from hetu_61a7_b200.models.nlp import BertConfig, BertForPreTraining
    cfg = (BertConfig.large if name == 'bert_large' else BertConfig.base)(batch_size=32, seq_len=128, attention_probs_dropout_prob=0.0)
    ids, tt, am, mlm, nsp = [ht.Variable(n) for n in ('ids', 'tt', 'am', 'mlm', 'nsp')]
    _, _, l1, l2 = BertForPreTraining(cfg)(ids, tt, am, mlm, nsp)
    loss = l1 + l2
    train = ht.optim.AdamOptimizer(1e-5).minimize(loss)
    ex = ht.Executor([loss, train], ctx=ctx)
    lab = rs.randint(0, cfg.vocab_size, (32, 128)).astype(np.float32)
    lab[rs.rand(32, 128) > 0.15] = -1
    fd = {ids: ht.array(rs.randint(0, cfg.vocab_size, (32, 128)).astype(np.float32), ctx), tt: ht.array(np.zeros((32, 128), np.float32), ctx),
          am: ht.array(np.ones((32, 128), np.float32), ctx), mlm: ht.array(lab, ctx), nsp: ht.array(rs.randint(0, 2, 32).astype(np.float32), ctx)}
    units = 32 * 128
for i in range(steps):
    torch.cuda.synchronize(); t0 = time.time()
    l = float(ex.run(feed_dict=fd)[0].asnumpy())
    torch.cuda.synchronize()
    print('step %d loss %.4f %.2f ms  (%.0f units/s)' % (i, l, (time.time() - t0) * 1e3, units / (time.time() - t0)), flush=True)
print('peak device memory %.2f GB (allocated), %.2f GB (reserved)' % (torch.cuda.max_memory_allocated() / 2**30, torch.cuda.max_memory_reserved() / 2**30))
