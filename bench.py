#!/usr/bin/env python
"""Headline benchmark (BASELINE.json): Wide&Deep-Criteo training samples/s (primary metric)
and MoE-layer training tokens/s (reported under "moe"), device-timed, max over ranks.

    python bench.py --gpus N --steps K --warmup W          # N == 1
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...

Config = the reference's own scripts: examples/ctr/run_hetu.py + models/wdl_criteo.py
(13 dense + 26 sparse fields, 33 762 577 x 128 fp32 embedding table, MLP 13-256-256-256,
batch 128 per GPU, SGD lr 0.01) and examples/moe/test_moe_top.py (16 x 1024 tokens,
d = h = 2048, 2 experts per GPU, top-2).  Synthetic data, random-init weights.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default='both', choices=['wdl', 'moe', 'both'])
    ap.add_argument('--batch', type=int, default=128, help='WDL batch per GPU (reference: 128)')
    ap.add_argument('--vocab', type=int, default=33762577)
    ap.add_argument('--moe-steps', type=int, default=0, help='0 = same as --steps')
    ap.add_argument('--collectives', default=None, choices=[None, 'symm', 'nccl'])
    ap.add_argument('--embedding', default='hbm', choices=['hbm', 'het'],
                    help="W&D table: 'hbm' = HBM-resident (sharded over NVLink for N > 1); 'het' = BASELINE config 3 as written: "
                         "host-DRAM parameter server + HET cache (GPU tier) + dense all-reduce")
    ap.add_argument('--profile', default=None, help='write a per-kernel time table of 50 W&D steps (torch.profiler) here')
    return ap.parse_args()


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
         'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu_index = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '--query-gpu=' + self.Q, '--format=csv,noheader,nounits', '-lms', '20',
                 '-i', str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
            t0 = time.time()
            while not self.rows and time.time() - t0 < 5.0:      # nvidia-smi needs a moment to start: the timed
                time.sleep(0.01)                                  # region must not begin before it is sampling
        except Exception:
            self.proc = None

    def mark(self):
        """Timed region starts now: rows sampled before (idle clocks) only serve as a fallback."""
        self.pre_rows, self.rows = self.rows, []

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        n0 = len(self.rows)
        t0 = time.time()
        while len(self.rows) <= n0 and time.time() - t0 < 1.0:   # one more sample after the last timed step
            time.sleep(0.01)
        self.proc.terminate()
        if not self.rows:
            self.rows = getattr(self, 'pre_rows', [])
        sm = sorted(float(r[1]) for r in self.rows if len(r) > 2 and r[1].replace('.', '').isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) > 2 and r[2].replace('.', '').isdigit()]
        reasons = set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            for i, n in enumerate(names):
                if len(r) > 5 + i and r[5 + i].lower().startswith('active'):
                    reasons.add(n)
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': sorted(reasons), 'samples': len(self.rows)}


def reference_arm(args):
    """Replace this process by baseline/ref_arm.py: the unmodified reference (copy of /root/reference under
    baseline/_ref, built by baseline/build_reference.sh) driven through its own API.  The new process has neither
    this repository's root nor its packages on sys.path, so nothing of hetu_61a7_b200 / libhetu_*.so is loaded."""
    ref = os.path.join(ROOT, 'baseline', '_ref')
    lib = os.path.join(ref, 'build', 'lib', 'libc_runtime_api.so')
    if not os.path.exists(lib) and int(os.environ.get('LOCAL_RANK', '0')) == 0:
        if os.path.isdir(os.path.join(ref, 'src')) or os.path.isdir('/root/reference/src'):
            subprocess.run(['bash', os.path.join(ROOT, 'baseline', 'build_reference.sh')],
                           stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    t0 = time.time()
    while not os.path.exists(lib) and int(os.environ.get('LOCAL_RANK', '0')) != 0 and time.time() - t0 < 900:
        time.sleep(1.0)
    if not os.path.exists(lib):
        if int(os.environ.get('RANK', '0')) == 0:
            print(json.dumps({'impl': 'reference', 'unavailable': 'baseline/_ref is not present / did not build on '
                              'this machine (baseline/build_reference.sh needs /root/reference or baseline/_ref/src)'}))
        return 0
    env = dict(os.environ)
    env.pop('PYTHONPATH', None)
    argv = [sys.executable, os.path.join(ROOT, 'baseline', 'ref_arm.py'), '--gpus', str(args.gpus),
            '--steps', str(args.steps), '--warmup', str(args.warmup), '--workload', args.workload,
            '--batch', str(args.batch), '--moe-steps', str(args.moe_steps)]
    sys.stdout.flush()
    os.execve(sys.executable, argv, env)


def timed(fn, steps, torch, device, comm, min_seconds=0.5, max_blocks=40, finish=None):
    """Blocks of EXACTLY `steps` calls of fn, each bracketed by barrier + synchronize on both sides and timed with CUDA
    events, MAX over ranks.  Blocks repeat until at least three have run and `min_seconds` of timed work has
    accumulated (short steps would otherwise be timed over a few milliseconds only); the median block is reported.
    -> (ms/step of the median block, {'blocks': n, 'min': .., 'max': ..})"""
    from hetu_61a7_b200.parallel.comm import device_max_ms
    per_step, total = [], 0.0
    while True:
        comm.barrier()
        torch.cuda.synchronize(device)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for i in range(steps):
            fn(i)
        if finish is not None:
            finish()
        e.record()
        torch.cuda.synchronize(device)
        comm.barrier()
        ms = device_max_ms(s.elapsed_time(e))           # identical on every rank: the loop exit is collective
        per_step.append(ms / steps)
        total += ms
        if (len(per_step) >= 3 and total >= min_seconds * 1e3) or len(per_step) >= max_blocks:
            break
    srt = sorted(per_step)
    return srt[len(srt) // 2], {'blocks': len(srt), 'min_ms_per_step': srt[0], 'max_ms_per_step': srt[-1]}


def start_parameter_server(world):
    """--embedding het: a scheduler and one server process on this node (what `heturun -s 1` starts), workers = the ranks."""
    port = int(os.environ.get('MASTER_PORT', '29500')) + 211
    env = {'DMLC_PS_ROOT_URI': '127.0.0.1', 'DMLC_PS_ROOT_PORT': str(port), 'DMLC_NUM_WORKER': str(world), 'DMLC_NUM_SERVER': '1'}
    os.environ.update(env)
    procs = []
    if int(os.environ.get('LOCAL_RANK', '0')) == 0:
        e = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get('PYTHONPATH', ''))
        for role in ('start_sched', 'start_server'):
            procs.append(subprocess.Popen([sys.executable, '-c', 'import hetu_61a7_b200.launcher as l; l.%s()' % role], env=e))
            time.sleep(0.5)
    return procs


def bench_wdl(args, ht, torch, np, comm, device):
    from hetu_61a7_b200 import _lib
    from hetu_61a7_b200.models import ctr
    world, rank = comm.nrank, comm.rank
    ctx = ht.gpu(device.index)
    B = args.batch
    npool = 32
    pool = []
    for i in range(npool):
        d, s, l = ht.data.criteo_synthetic(B, vocab=args.vocab, seed=1000 * rank + i)
        pool.append((d, s.astype(np.int64), l))
    # inputs are dataloader nodes over a pinned host dataset (public API: ht.dataloader_op); the
    # device-timed loop below overrides them with HBM-resident feeds
    dense = ht.dataloader_op([ht.Dataloader(np.concatenate([p[0] for p in pool]), B, 'train')])
    sparse = ht.dataloader_op([ht.Dataloader(np.concatenate([p[1] for p in pool]), B, 'train', dtype=np.int64)])
    y_ = ht.dataloader_op([ht.Dataloader(np.concatenate([p[2] for p in pool]), B, 'train')])
    sharded = world > 1
    loss, pred, _, train_op = ctr.wdl_criteo(dense, sparse, y_, feature_dimension=args.vocab, ctx=ctx)
    kw = dict(ctx=ctx, seed=123, shard_dataloaders=False)    # every rank generates its own shard
    if args.embedding == 'het':
        from hetu_61a7_b200.graph import executor as _ex
        _ex.worker_init()
        kw.update(comm_mode='Hybrid', cstable_policy='LFUOpt', cache_bound=3, bsp=-1)      # examples/ctr/tests/hybrid_wdl_criteo.sh:7
    elif world > 1:
        kw['comm_mode'] = 'Hybrid'
        if args.collectives:
            kw['use_nccl_collectives'] = args.collectives == 'nccl'
    ex = ht.Executor({'train': [loss, pred, train_op]}, **kw)
    # --- device-timed: inputs resident in HBM, graph replays back to back ---------------
    # eight batches packed [dense | ids | labels] in HBM; the fed arrays are views of ONE buffer, so a step's inputs change
    # with one device-to-device copy of 33 KB (three separate copies cost ~4 us of a 43 us step)
    def _pack(d, s_, l):
        parts = [np.ascontiguousarray(d, dtype=np.float32).view(np.uint8).reshape(-1),
                 np.ascontiguousarray(s_, dtype=np.int64).view(np.uint8).reshape(-1),
                 np.ascontiguousarray(l, dtype=np.float32).view(np.uint8).reshape(-1)]
        offs, total = [], 0
        for p_ in parts:
            offs.append(total)
            total += (p_.size + 255) // 256 * 256
        flat = np.zeros(total, dtype=np.uint8)
        for o, p_ in zip(offs, parts):
            flat[o:o + p_.size] = p_
        return flat, offs, [p_.size for p_ in parts]

    packed = [_pack(*pool[i]) for i in range(8)]
    dev_pool = [torch.from_numpy(f).to(device) for f, _, _ in packed]
    offs, sizes = packed[0][1], packed[0][2]
    in_buf = dev_pool[0].clone()
    d_buf = ht.NDArray(in_buf[offs[0]:offs[0] + sizes[0]].view(torch.float32).view(tuple(pool[0][0].shape)))
    s_buf = ht.NDArray(in_buf[offs[1]:offs[1] + sizes[1]].view(torch.int64).view(tuple(pool[0][1].shape)))
    l_buf = ht.NDArray(in_buf[offs[2]:offs[2] + sizes[2]].view(torch.float32).view(tuple(pool[0][2].shape)))

    launches0 = _lib.launch_count
    total_steps = [0]

    def step_dev(i):
        total_steps[0] += 1
        in_buf.copy_(dev_pool[i % len(dev_pool)], non_blocking=True)
        ex.run('train', feed_dict={dense: d_buf, sparse: s_buf, y_: l_buf})

    for i in range(max(args.warmup, 3)):
        step_dev(i)
    sub = ex.subexecutor['train']
    sampler = ClockSampler(device.index)
    sampler.start()
    for i in range(3):
        step_dev(i)
    sampler.mark()
    ms_dev, spread_dev = timed(step_dev, args.steps, torch, device, comm)
    if args.profile:                                # diagnostics only: kernel durations inside the replayed graph
        from torch.profiler import profile, ProfilerActivity
        comm.barrier()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for i in range(50):
                step_dev(i)
            torch.cuda.synchronize(device)
        if rank == 0:
            with open(args.profile, 'w') as fw:
                fw.write('W&D step, %d GPU(s), 50 steps, torch.profiler (CUPTI) kernel times\n' % world)
                fw.write(prof.key_averages().table(sort_by='cuda_time_total', row_limit=25, max_name_column_width=90))
        comm.barrier()
    if os.environ.get('HETU_WDL_STAMPS'):           # diagnostics: phase time stamps of the last fused tower launch
        import ctypes
        out = (ctypes.c_ulonglong * 16)()
        _lib.call('wdl_tower_stamps', out)
        names = ['P0', 'P1', 'P2', 'P3', 'head', 'P5', 'P6a', 'P6b', 'tail']
        row = ' '.join('%s %.2f' % (n, (out[i + 1] - out[i]) / 1e3) for i, n in enumerate(names) if out[i + 1] and out[i])
        rel = ' '.join('%s %.2f' % (n, (out[k] - out[0]) / 1e3) for n, k in (('gather0', 10), ('gather1', 11), ('sparse0', 12), ('sparse1', 13)))
        print('[rank %d] tower phases (us): %s | relative to tower start: %s' % (rank, row, rel), flush=True)
    # --- end to end: numpy batch -> pinned -> H2D, loss D2H every step ---------------------
    h2d = sum(a.nbytes for a in pool[0])
    last = {}

    from collections import deque
    inflight = deque()
    depth = max(1, int(os.environ.get('HETU_BENCH_READBACK_LAG', '1')))

    def step_e2e(i):
        out = ex.run('train')                   # batch i: its H2D copy runs beside the replay of step i-1
        inflight.append(out[0].asnumpy_async())  # loss i: D2H queued behind step i ...
        if len(inflight) > depth:               # ... and read one step later (a lag of 2 measured the same), so the host never waits on the newest step
            last['loss'] = inflight.popleft().get()

    def finish_e2e():
        while inflight:
            last['loss'] = inflight.popleft().get()

    for i in range(max(args.warmup, 3)):
        step_e2e(i)
    ms_e2e, spread_e2e = timed(step_e2e, args.steps, torch, device, comm, finish=finish_e2e)
    clocks = sampler.stop()
    launches = getattr(sub, 'graph_launches', 0) * args.steps if sub._graph is not None else 0
    cache_perf = None
    if args.embedding == 'het':
        caches = [getattr(n, 'cache', None) for n in ex.param_nodes if getattr(n, 'cache', None) is not None]
        cache_perf = caches[0].perf() if caches else None
        launches = (_lib.launch_count - launches0) * args.steps // max(total_steps[0], 1)
    return {
        'ms_dev': ms_dev, 'ms_e2e': ms_e2e, 'clocks': clocks, 'h2d': h2d, 'd2h': 4,
        'global_batch': B * world, 'launches': launches, 'graph': sub._graph is not None,
        'loss': float(np.asarray(last['loss']).reshape(-1)[0]),
        'launches_per_step': getattr(sub, 'graph_launches', 0), 'spread': spread_dev, 'spread_e2e': spread_e2e,
        'fused_wdl': getattr(ex, 'fused_wdl', 0), 'cache_perf': cache_perf,
    }


def bench_moe(args, ht, torch, np, comm, device, steps):
    from hetu_61a7_b200.models import moe
    world, rank = comm.nrank, comm.rank
    ctx = ht.gpu(device.index)
    bs, ntok, dim, hid, nle, top = 16, 1024, 2048, 2048, 2, 2
    # the step's input is a dataloader node: batches are views of a pinned host dataset, the executor
    # copies batch i+1 to HBM on its copy stream while step i computes (public API: ht.dataloader_op)
    rng = np.random.RandomState(rank)
    nbatch = 4
    data = rng.randn(nbatch * bs, ntok, dim).astype(np.float32)
    x = ht.dataloader_op([ht.Dataloader(data, bs, 'train')])
    t = ht.Variable(name='t')
    loss, _ = moe.moe_model(x, t, batch_size=bs, num_tokens=ntok, model_dim=dim, hidden_size=hid,
                            num_local_experts=nle, world=world, rank=rank, top=top)
    train = ht.optim.SGDOptimizer(learning_rate=0.125 / 1000).minimize(loss)
    kw = dict(ctx=ctx, seed=7, shard_dataloaders=False)      # every rank generates its own shard
    if world > 1:
        kw['comm_mode'] = 'AllReduce'
        if args.collectives:
            kw['use_nccl_collectives'] = args.collectives == 'nccl'
    ex = ht.Executor({'train': [loss, train]}, **kw)
    tv = np.zeros((bs,), dtype=np.float32)
    last = {}
    # (a) the reference's own protocol (examples/moe/test_moe_top.py:61-97): x lives on the GPU,
    #     every step = executor.run + loss.asnumpy()
    x_dev = ht.NDArray(torch.from_numpy(data[:bs]).to(device))
    t_dev = ht.array(tv, ctx)

    def step_dev(i):
        out = ex.run('train', feed_dict={x: x_dev, t: t_dev})
        last['loss'] = out[0].asnumpy()

    for i in range(max(3, min(args.warmup, 4))):
        step_dev(i)
    ms_dev, spread_dev = timed(step_dev, steps, torch, device, comm)

    # (b) end to end: every step's input batch comes from pinned host memory (H2D inside the timed
    #     region, overlapped with the previous step's compute) and the loss is read back every step
    def step(i):
        out = ex.run('train', feed_dict={t: tv})
        fut = out[0].asnumpy_async()
        if last.get('pending') is not None:
            last['loss'] = last['pending'].get()
        last['pending'] = fut

    def finish():
        if last.get('pending') is not None:
            last['loss'] = last.pop('pending').get()

    for i in range(3):
        step(i)
    ms, spread_e2e = timed(step, steps, torch, device, comm, finish=finish)
    sub = ex.subexecutor['train']
    return {'ms_e2e': ms, 'ms_dev': ms_dev, 'tokens': bs * ntok * world, 'h2d': bs * ntok * dim * 4 + tv.nbytes, 'd2h': 4,
            'loss': float(np.asarray(last['loss']).reshape(-1)[0]),
            'launches_per_step': getattr(sub, 'graph_launches', 0), 'graph': sub._graph is not None,
            'spread': spread_dev, 'spread_e2e': spread_e2e}


def main():
    args = parse()
    if args.impl == 'reference':
        return reference_arm(args)
    import numpy as np
    import torch
    if not torch.cuda.is_available():
        print(json.dumps({'metric': 'wdl_criteo_samples_per_s', 'value': None, 'error': 'no GPU visible'}))
        return 1
    import __graft_entry__ as ge
    if not os.path.exists(os.path.join(ge.LIB, 'libhetu_b200.so')):
        if int(os.environ.get('LOCAL_RANK', '0')) == 0:
            ge.build()
    ps_procs = start_parameter_server(int(os.environ.get('WORLD_SIZE', '1'))) if args.embedding == 'het' else []
    import hetu_61a7_b200 as ht
    from hetu_61a7_b200.parallel.comm import world as get_world
    comm = ht.wrapped_mpi_nccl_init()
    device = torch.device('cuda', torch.cuda.current_device())
    world = comm.nrank
    assert world == args.gpus or world == 1, 'launched with %d ranks but --gpus %d' % (world, args.gpus)
    out = {'metric': 'wdl_criteo_samples_per_s', 'unit': 'samples/s', 'n_gpus': world, 'steps': args.steps,
           'warmup': max(args.warmup, 3), 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
           'dtype': 'fp32 storage, tf32 tensor-core GEMM (reference precision family: fp32)',
           'data': 'synthetic (Criteo-shaped zipf ids; random-init weights)', 'impl': 'ours'}
    if args.workload in ('wdl', 'both'):
        w = bench_wdl(args, ht, torch, np, comm, device)
        out.update({
            'value': w['global_batch'] / (w['ms_dev'] / 1e3), 'ms_per_step': w['ms_dev'],
            'e2e': {'value': w['global_batch'] / (w['ms_e2e'] / 1e3), 'unit': 'samples/s',
                    'ms_per_step': w['ms_e2e'], 'h2d_bytes_per_step': w['h2d'], 'd2h_bytes_per_step': w['d2h']},
            'gpu_launches': w['launches'], 'clocks': w['clocks'],
            'config': {'model': 'Wide&Deep (13 dense + 26 sparse, %d x 128 fp32 table, MLP 13-256-256-256)' % args.vocab,
                       'global_batch': w['global_batch'], 'seq_len': None,
                       'parallelism': 'dp%d' % world if world > 1 else 'single',
                       'embedding': ('host-DRAM parameter server + HET cache in HBM (LFUOpt, bound 3), dense all-reduce'
                                     if args.embedding == 'het' else
                                     'HBM-resident' + (' sharded over NVLink' if world > 1 else '')),
                       'het_cache': w.get('cache_perf'),
                       'optimizer': 'SGD lr 0.01', 'cuda_graph': w['graph'],
                       'kernel_launches_per_step': w['launches_per_step'],
                       'l2_policy': 'working set (17.3 GB embedding table, random rows) exceeds the 126 MB L2',
                       'final_loss': w['loss'], 'fused_step': bool(w['fused_wdl']),
                       'timing': dict(w['spread'], rule='median of blocks of `steps` steps, >= 3 blocks and >= 0.5 s'),
                       'e2e_timing': dict(w['spread_e2e'], loss_readback='async D2H to pinned memory every step, '
                                          'consumed one step later')},
        })
    if args.workload in ('moe', 'both'):
        steps = args.moe_steps or args.steps
        m = bench_moe(args, ht, torch, np, comm, device, steps)
        moe_out = {'metric': 'moe_tokens_per_s', 'value': m['tokens'] / (m['ms_dev'] / 1e3), 'unit': 'tokens/s',
                   'ms_per_step': m['ms_dev'], 'steps': steps,
                   'e2e': {'value': m['tokens'] / (m['ms_e2e'] / 1e3), 'unit': 'tokens/s', 'ms_per_step': m['ms_e2e'],
                           'h2d_bytes_per_step': m['h2d'], 'd2h_bytes_per_step': m['d2h']},
                   'final_loss': m['loss'], 'cuda_graph': m['graph'], 'timing': m['spread'], 'e2e_timing': m['spread_e2e'],
                   'kernel_launches_per_step': m['launches_per_step'],
                   'config': {'model': 'MoE layer top-2, 16x1024 tokens/GPU, d=h=2048, 2 experts/GPU',
                              'global_tokens': m['tokens'], 'parallelism': 'ep%d' % world}}
        if args.workload == 'moe':
            out.update({'metric': 'moe_tokens_per_s', 'value': moe_out['value'], 'unit': 'tokens/s',
                        'ms_per_step': m['ms_dev'], 'config': moe_out['config'], 'e2e': moe_out['e2e'],
                        'gpu_launches': m['launches_per_step'] * steps})
        else:
            out['moe'] = moe_out
    if comm.rank == 0:
        print(json.dumps(out))
        sys.stdout.flush()
    if args.embedding == 'het':
        comm.barrier()
        for p_ in ps_procs:
            p_.terminate()
        os._exit(0)
    return 0


if __name__ == '__main__':
    sys.exit(main())
