"""Single-process multi-GPU communicator (late-collected: needs torch.cuda.nccl with several devices in one process)."""
import pytest


@pytest.mark.gpu
def test_single_process_multi_gpu_communicator():
    """reference communicator/nccl_comm.py: one process, one array per local GPU."""
    import numpy as np
    import torch
    from hetu.communicator.nccl_comm import NCCL_Communicator
    n = min(torch.cuda.device_count(), 4)
    comm = NCCL_Communicator(list(range(n)))
    xs = [torch.full((8, 4), float(d + 1), device='cuda:%d' % d) for d in range(n)]
    try:
        comm.all_reduce(xs)
    except (RuntimeError, AttributeError) as e:                 # torch.cuda.nccl is an optional, legacy torch module
        pytest.skip('torch.cuda.nccl not usable in this process: %s' % e)
    for x in xs:
        assert torch.all(x == sum(range(1, n + 1)))
    ins = [torch.arange(n * 2 * 3, dtype=torch.float32, device='cuda:%d' % d).reshape(n * 2, 3) + 100 * d for d in range(n)]
    outs = [torch.empty_like(t) for t in ins]
    comm.all_to_all(ins, outs)
    for j in range(n):
        for i in range(n):
            np.testing.assert_array_equal(outs[j][i * 2:(i + 1) * 2].cpu().numpy(), ins[i][j * 2:(j + 1) * 2].cpu().numpy())
    g_out = [torch.empty((n * 8, 4), device='cuda:%d' % d) for d in range(n)]
    try:
        comm.all_gather(xs, g_out)
    except (RuntimeError, AttributeError) as e:
        pytest.skip('torch.cuda.nccl all_gather not usable: %s' % e)
    assert g_out[0].shape[0] == n * 8
