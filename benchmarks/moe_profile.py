"""Per-kernel time table of the MoE training step (torch.profiler / CUPTI), one or several GPUs.

    python benchmarks/moe_profile.py out.txt                       # 1 GPU
    torchrun --nproc-per-node 2 ... benchmarks/moe_profile.py out.txt

Diagnostics only (profiler overhead is inside the numbers); bench.py gives the timed figures.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hetu_61a7_b200 as ht                                            # noqa: E402
from hetu_61a7_b200.models import moe                                   # noqa: E402


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/moe_profile.txt'
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    torch.cuda.set_device(local)
    bs, ntok, dim, hid = 16, 1024, 2048, 2048
    x, t = ht.Variable('x'), ht.Variable('t')
    loss, _ = moe.moe_model(x, t, batch_size=bs, num_tokens=ntok, model_dim=dim, hidden_size=hid,
                            num_local_experts=2, world=world, rank=rank, top=2)
    train = ht.optim.SGDOptimizer(learning_rate=0.125 / 1000).minimize(loss)
    kw = dict(ctx=ht.gpu(local), seed=7)
    if world > 1:
        kw['comm_mode'] = 'AllReduce'
    ex = ht.Executor({'train': [loss, train]}, **kw)
    xv = ht.NDArray(torch.randn(bs, ntok, dim, device='cuda:%d' % local))
    tv = ht.array(np.zeros((bs,), np.float32), ht.gpu(local))
    for _ in range(5):
        ex.run('train', feed_dict={x: xv, t: tv})
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity
    steps = 10
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(steps):
            ex.run('train', feed_dict={x: xv, t: tv})
        torch.cuda.synchronize()
    if rank == 0:
        rows = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)
        with open(out_path, 'w') as f:
            f.write('MoE step, %d GPU(s), %d steps, HETU_MOE_OVERLAP=%s\n' % (world, steps, os.environ.get('HETU_MOE_OVERLAP', '1')))
            tot = sum(e.device_time_total for e in rows)
            f.write('sum of kernel time per step: %.1f us\n' % (tot / steps))
            for e in rows[:40]:
                f.write('%9.1f us/step  x%-4.1f  %s\n' % (e.device_time_total / steps, e.count / steps, e.key[:110]))
    if world > 1:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.barrier()


if __name__ == '__main__':
    main()
