"""Probe: does torch's symmetric-memory allocator give a multicast (NVLS) mapping on this box?
torchrun --nproc-per-node N benchmarks/probe_symm_mem.py"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ['RANK'])
torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
dist.init_process_group('nccl')
t = symm_mem.empty(1 << 20, dtype=torch.float32, device='cuda')
h = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
print('rank', rank, 'world', h.world_size, 'buffer_ptrs', [hex(p) for p in h.buffer_ptrs][:3], 'multicast_ptr', hex(h.multicast_ptr),
      'signal_pad', hex(h.signal_pad_ptrs[0]), flush=True)
t.fill_(rank + 1)
dist.barrier()
torch.cuda.synchronize()
dist.destroy_process_group()
