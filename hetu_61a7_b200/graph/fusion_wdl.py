"""Whole-step fusion of the Wide&Deep training graph (BASELINE config 3).

The pass recognises the graph of the reference's examples/ctr/models/wdl_criteo.py:8-42 -

    sparse = reshape(embedding_lookup(E, ids));  h1 = relu(x W1);  h2 = relu(h1 W2);  y3 = h2 W3
    y = sigmoid(concat(sparse, y3) W4);  loss = mean(bce(y, y_));  SGD.minimize(loss)

- after autodiff and the optimizer hooks have run, and replaces the ~45 forward / backward / update nodes by
three kernels (csrc/wdl_tower.cu): a gather that writes the looked-up rows straight into the concat buffer,
ONE cluster kernel for tower forward + loss + tower backward on tcgen05 (batch 128 = one MMA tile), and a
rank-one SGD scatter into the table; the dense SGD (or, under data parallelism, the fused all-reduce + SGD over
the symmetric heap) follows as before.  The reference executes every node as its own kernel launch from Python
(python/hetu/gpu_ops/executor.py:1000-1056).

The rewrite keeps the user's nodes: the loss, prediction and optimizer nodes stay in the graph (they are what the
user evaluates) but read from the fused node.  Anything that does not match exactly - other optimizers, l2
regularisation, another batch size, a PS-resident table, extra consumers of an intermediate - leaves the graph
untouched.  ``HETU_FUSE_WDL=0`` switches the pass off.
"""
from __future__ import annotations

import ctypes
import os

import torch

from .. import _lib as L

_IDX_CODE = {torch.float32: 0, torch.int32: 1, torch.int64: 2}
TOWER_BATCH = 128
TOWER_HIDDEN = 256


def _is(node, name):
    return node is not None and node.op_type == name


def _match(opt_node):
    """Returns the dict of matched nodes or None."""
    from ..optimizer import OPT_SGD
    from ..ops.variable import PlaceholderOp
    opt = opt_node.optimizer
    if opt.kind != OPT_SGD or float(opt.l2reg) != 0.0:
        return None
    loss = opt.loss
    if isinstance(loss, (list, tuple)) or not _is(loss, 'ReduceMeanOp') or list(loss.axes or []) != [0] or loss.keepdims:
        return None
    bce = loss.inputs[0]
    if not _is(bce, 'BinaryCrossEntropyOp'):
        return None
    y, labels = bce.inputs
    if not _is(y, 'SigmoidOp'):
        return None
    mm4 = y.inputs[0]

    def plain_mm(n):
        return _is(n, 'MatMulOp') and len(n.inputs) == 2 and not n.matmul_attr_trans_A and not n.matmul_attr_trans_B \
            and getattr(n, 'act', 0) == 0

    if not plain_mm(mm4):
        return None
    y4, W4 = mm4.inputs
    if not _is(y4, 'ConcatOp') or y4.axis != 1:
        return None
    sparse, y3 = y4.inputs
    if not _is(sparse, 'ArrayReshapeOp') or not _is(sparse.inputs[0], 'EmbeddingLookUp'):
        return None
    lookup = sparse.inputs[0]
    E, ids = lookup.inputs
    if not plain_mm(y3):
        return None
    r2, W3 = y3.inputs
    if not _is(r2, 'ReluOp') or not plain_mm(r2.inputs[0]):
        return None
    r1, W2 = r2.inputs[0].inputs
    if not _is(r1, 'ReluOp') or not plain_mm(r1.inputs[0]):
        return None
    x, W1 = r1.inputs[0].inputs
    params = [E, W1, W2, W3, W4]
    if any(not isinstance(p, PlaceholderOp) or not p.trainable for p in params):
        return None
    if set(map(id, opt.params)) != set(map(id, params)):
        return None
    if getattr(E, 'ps_resident', False):
        return None
    shp = [tuple(getattr(p, 'shape', None) or ()) for p in params]
    if len(shp[0]) != 2 or len(shp[1]) != 2:
        return None
    nd, hid = shp[1]
    width = shp[0][1]
    if hid != TOWER_HIDDEN or shp[2] != (hid, hid) or shp[3] != (hid, hid) or not (1 <= nd <= 16):
        return None
    if len(shp[4]) != 2 or shp[4][1] != 1 or (shp[4][0] - hid) <= 0 or (shp[4][0] - hid) % width != 0:
        return None
    emb_cols = shp[4][0] - hid
    if emb_cols % 32 != 0 or width % 4 != 0:
        return None
    interior = [bce, mm4, y4, sparse, lookup, y3, r2, r2.inputs[0], r1, r1.inputs[0]]
    return dict(loss=loss, pred=y, labels=labels, x=x, ids=ids, E=E, W1=W1, W2=W2, W3=W3, W4=W4, nd=nd, width=width,
                emb_cols=emb_cols, fields=emb_cols // width, interior=interior)


_OP_CLASS = []


def _make_op_class():
    """The op class is built on first use: graph.node imports this package."""
    if _OP_CLASS:
        return _OP_CLASS[0]
    from .node import Op

    class _WdlTowerStepOp(Op):
        def __init__(self, m, ctx=None):
            super().__init__(_WdlTowerStepOp, [m['x'], m['ids'], m['labels'], m['E'], m['W1'], m['W2'], m['W3'], m['W4']], ctx)
            self.m = m
            self._bufs = None
            self._dp = None
            self.last = None
            self.fused_launches = 0

        def infer_shape(self, input_shapes):
            return None

        def gradient(self, g):
            raise NotImplementedError('the fused Wide&Deep step already contains its backward pass')

        # ---- reference semantics in plain torch (CPU, other batch sizes) ------------------------------------
        def _torch_step(self, x, ids, labels, E, W1, W2, W3, W4):
            m = self.m
            idx = ids.long() if ids.dtype != torch.int64 else ids
            rows = E.lookup(idx) if hasattr(E, 'lookup') else E[idx]
            B = x.shape[0]
            h1 = torch.relu(x @ W1)
            h2 = torch.relu(h1 @ W2)
            y4 = torch.cat([rows.reshape(B, m['emb_cols']), h2 @ W3], 1)
            pr = torch.sigmoid(y4 @ W4)
            lab = labels.reshape(B, 1)
            loss = (-lab * torch.log(pr) - (1 - lab) * torch.log(1 - pr)).mean(0)
            dz = ((1.0 / B) * (-lab / pr + (1 - lab) / (1 - pr)) * pr * (1 - pr)).reshape(B)
            w4 = W4.reshape(-1)
            dW4 = (y4.t() @ dz).reshape(-1, 1)
            dy3 = dz[:, None] * w4[m['emb_cols']:][None, :]
            dW3 = h2.t() @ dy3
            dh2 = (dy3 @ W3.t()) * (h2 > 0)
            dW2 = h1.t() @ dh2
            dh1 = (dh2 @ W2.t()) * (h1 > 0)
            dW1 = x.t() @ dh1
            self.last = dict(ids=idx, dz=dz, grads=(dW1, dW2, dW3, dW4), W4=W4, E=E, torch_path=True)
            return (loss, pr)

        def compute(self, input_vals, output_val=None, stream_handle=None):
            x, ids, labels, E, W1, W2, W3, W4 = input_vals
            m = self.m
            if x.device.type == 'meta':
                return (torch.empty(1, device='meta'), torch.empty((x.shape[0], 1), device='meta'))
            dev = x.device
            on_gpu = dev.type == 'cuda' and L.cuda_lib() is not None and x.shape[0] == TOWER_BATCH \
                and x.dtype == torch.float32 and ids.dtype in _IDX_CODE
            if not on_gpu:
                return self._torch_step(x, ids, labels, E, W1, W2, W3, W4)
            B, hid, pitch = TOWER_BATCH, TOWER_HIDDEN, m['emb_cols'] + TOWER_HIDDEN
            if self._bufs is None:
                f = dict(dtype=torch.float32, device=dev)
                self._bufs = dict(y4=torch.empty((B, pitch), **f), work=torch.empty(3 * B * hid, **f),
                                  loss=torch.empty(1, **f), pred=torch.empty((B, 1), **f), dz=torch.empty(B, **f),
                                  zpf=torch.empty((B, m['fields']), **f),
                                  dW1=torch.empty((m['nd'], hid), **f), dW2=torch.empty((hid, hid), **f),
                                  dW3=torch.empty((hid, hid), **f), dW4=torch.empty((pitch, 1), **f))
            b = self._bufs
            if hasattr(E, 'lookup'):                       # ShardedTable: peer base pointers over the symmetric heap
                bases, nshard, vocab = E.buf.ptrs, E.world, E.vocab
            else:
                bases, nshard, vocab = (ctypes.c_void_p * 1)(E.data_ptr()), 1, E.shape[0]
            x, ids, labels = x.contiguous(), ids.contiguous(), labels.contiguous()
            n = ids.numel()
            st = L.stream_ptr(dev)
            dp = self._dp_state(dev, pitch)
            # programmatic dependent launches: gather || first tower layers, sparse SGD || tower backward.  Not under
            # BSP, where cross-GPU barrier kernels sit between the tower and the table update.
            pdl = os.environ.get('HETU_WDL_PDL', '1') == '1' and getattr(E, 'bsp', -1) != 0
            if pdl and 'dz_flag' not in b:
                b['dz_flag'] = torch.zeros(1, dtype=torch.int32, device=dev)
                b['dz_expect'] = torch.zeros(1, dtype=torch.int32, device=dev)
            L.call('gather_rows_pitched', bases, L.c_int(nshard), L.p(ids), L.c_int(_IDX_CODE[ids.dtype]), L.p(b['y4']),
                   L.c_sz(n), L.c_int(m['width']), L.c_int(m['fields']), L.c_int(pitch), L.c_sz(vocab), L.p(W4),
                   L.p(b['zpf']), L.p(b['w4_snap'] if dp is not None else None),
                   L.p(b['dz_flag'] if pdl else None), L.p(b['dz_expect'] if pdl else None), st)
            if dp is None:
                tail = (None, None, L.c_int(1), L.c_int(0), L.p(None), L.p(None), L.p(None))
            else:
                tail = (dp['arenas'], dp['flags'], L.c_int(dp['world']), L.c_int(dp['rank']), L.p(dp['epoch']),
                        L.p(dp['opt'].lr_t), L.p(dp['opt'].step_t))
            L.call('wdl_tower_step', L.p(x), L.p(b['y4']), L.p(b['zpf']), L.p(labels), L.p(W1), L.p(W2), L.p(W3), L.p(W4),
                   L.p(b['work']), L.p(b['loss']), L.p(b['pred']), L.p(b['dz']), L.p(b['dW1']), L.p(b['dW2']),
                   L.p(b['dW3']), L.p(b['dW4']), L.c_int(B), L.c_int(m['nd']), L.c_int(hid), L.c_int(m['emb_cols']),
                   L.c_int(m['width']), *tail, L.c_int(1 if pdl else 0), L.p(b['dz_flag'] if pdl else None),
                   ctypes.c_void_p((dp or {}).get('arena_mc', 0) or 0), ctypes.c_void_p((dp or {}).get('flags_mc', 0) or 0), st)
            self.fused_launches += 2
            self.last = dict(ids=ids, dz=b['dz'], grads=(b['dW1'], b['dW2'], b['dW3'], b['dW4']),
                             W4=b['w4_snap'] if dp is not None else W4, E=E, bases=bases, nshard=nshard, vocab=vocab,
                             torch_path=False, dense_done=dp is not None, weights=(W1, W2, W3, W4),
                             dz_flag=b['dz_flag'] if pdl else None, dz_expect=b['dz_expect'] if pdl else None)
            return (b['loss'], b['pred'])

        def _dp_state(self, dev, pitch):
            """In-kernel all-reduce + SGD (csrc/wdl_tower.cu): arena / flags / epoch, built on the first step.
            None when the dense update stays outside the kernel (no optimizer attached: kernel tests)."""
            if self._dp is not None:
                return self._dp or None
            opt_node = getattr(self, 'opt_node', None)
            if opt_node is None or os.environ.get('HETU_WDL_INKERNEL_SGD', '1') != '1':
                self._dp = False
                return None
            m = self.m
            opt = opt_node.optimizer
            total = m['nd'] * TOWER_HIDDEN + 2 * TOWER_HIDDEN * TOWER_HIDDEN + pitch
            self._bufs['w4_snap'] = torch.empty(m['emb_cols'], dtype=torch.float32, device=dev)
            if opt.step_t is None or opt.step_t.device != dev:
                opt.step_t = torch.zeros(1, dtype=torch.int64, device=dev) if opt.step_t is None else opt.step_t.to(dev)
                opt.lr_t = opt.lr_t.to(dev)
            fdp = opt_node.fused_dp
            if fdp is None:
                arena = torch.empty(2 * total, dtype=torch.float32, device=dev)
                self._dp = dict(world=1, rank=0, arenas=(ctypes.c_void_p * 1)(arena.data_ptr()), flags=None,
                                epoch=torch.zeros(1, dtype=torch.int32, device=dev), opt=opt, keep=arena)
                return self._dp
            eng = fdp.engine
            pos = {id(p): i for i, p in enumerate(opt.params)}
            if eng.world > 8 or sorted(fdp.ids) != sorted(pos[id(m[k])] for k in ('W1', 'W2', 'W3', 'W4')):
                self._dp = False
                return None
            arena_bytes = 2 * eng.world * total * 4
            mc = eng.mc_buffer(('wdl_arena', id(self)), arena_bytes + 256)       # arena + flag row, NVLS multicast mapped
            if mc is not None:
                self._dp = dict(world=eng.world, rank=eng.rank, arenas=mc.ptrs, flags=mc.peer_ptrs(arena_bytes),
                                arena_mc=mc.mc_ptr, flags_mc=mc.mc_ptr + arena_bytes,
                                arena_local=mc.local_ptr, flags_local=mc.local_ptr + arena_bytes,
                                epoch=torch.zeros(1, dtype=torch.int32, device=dev), opt=opt, keep=mc)
                return self._dp
            arena = eng.buffer(('wdl_arena', id(self)), arena_bytes)
            flags = eng.buffer(('wdl_flags', id(self)), 256)
            flags.tensor(dtype=torch.int32, numel=64).zero_()
            torch.cuda.synchronize(dev)
            eng.comm.barrier()
            self._dp = dict(world=eng.world, rank=eng.rank, arenas=arena.ptrs, flags=flags.ptrs,
                            arena_local=arena.local_ptr, flags_local=flags.local_ptr,
                            epoch=torch.zeros(1, dtype=torch.int32, device=dev), opt=opt, keep=(arena, flags))
            return self._dp

        # ---- the optimizer node's new compute -------------------------------------------------------------------
        def apply_update(self, opt_node):
            m, last = self.m, self.last
            opt = opt_node.optimizer
            E = last['E']
            pos = {id(p): i for i, p in enumerate(opt.params)}
            grads = [None] * len(opt.params)
            if not last.get('dense_done'):
                for name, g in zip(('W1', 'W2', 'W3', 'W4'), last['grads']):
                    grads[pos[id(m[name])]] = g
            bsp = getattr(E, 'bsp', -1) == 0 and hasattr(E, 'engine') and not last['torch_path']
            if bsp:
                E.engine.barrier()
            if last['torch_path']:
                from ..ndarray import IndexedSlices
                w4 = last['W4'].reshape(-1)[:m['emb_cols']].reshape(m['fields'], m['width'])
                vals = (last['dz'][:, None, None] * w4[None]).reshape(-1, m['width'])
                shape = tuple(getattr(m['E'], 'shape', None) or E.shape)
                grads[pos[id(m['E'])]] = IndexedSlices(last['ids'].reshape(-1), vals, shape)
            else:
                ids = last['ids']
                L.call('wdl_sparse_sgd_rank1', last['bases'], L.c_int(last['nshard']), L.p(ids),
                       L.c_int(_IDX_CODE[ids.dtype]), L.p(last['dz']), L.p(last['W4']), L.p(opt.lr_t), L.c_f32(1.0),
                       L.c_sz(ids.numel()), L.c_int(m['width']), L.c_int(m['fields']), L.c_sz(last['vocab']),
                       L.p(last.get('dz_flag')), L.p(last.get('dz_expect')), L.stream_ptr(ids.device))
            if bsp:
                E.engine.barrier()
            if last.get('dense_done'):
                dp = self._dp
                if dp and dp['world'] > 1:
                    # the tower kernel has pushed its gradient blocks to every rank (multimem.st) and raised its flag;
                    # sum the blocks that arrived here and apply SGD
                    W = last['weights']
                    L.call('wdl_dp_apply', ctypes.c_void_p(dp['arena_local']), ctypes.c_void_p(dp['flags_local']),
                           L.p(dp['epoch']), L.p(opt.lr_t), L.p(W[0]), L.p(W[1]), L.p(W[2]), L.p(W[3]),
                           L.c_int(m['nd']), L.c_int(m['emb_cols']), L.c_int(dp['world']), L.c_int(dp['rank']),
                           L.stream_ptr(W[0].device))
                return                                    # one GPU: SGD and the step counter were done inside the tower kernel
            if opt_node.fused_dp is not None:
                opt_node.fused_dp.step(opt, grads)
            else:
                opt.update(grads)

    _OP_CLASS.append(_WdlTowerStepOp)
    return _WdlTowerStepOp


def fuse_wide_deep(eval_nodes) -> int:
    """Rewrites every matching optimizer node reachable from ``eval_nodes``; returns how many were fused."""
    if os.environ.get('HETU_FUSE_WDL', '1') != '1':
        return 0
    from .autodiff import find_topo_sort
    from ..optimizer import OptimizerOp
    topo = find_topo_sort(list(eval_nodes))
    fused = 0
    for opt_node in [n for n in topo if isinstance(n, OptimizerOp)]:
        if getattr(opt_node, 'wdl_fused', None) is not None or opt_node.comm_mode == 'PS':
            continue
        # the gradients must reach the optimizer directly (the symmetric-heap DP step takes them as they are);
        # NCCL all-reduce / PS nodes in between mean another synchronisation plan: leave that graph alone
        if any(i is not None and i.op_type in ('AllReduceCommunicateOp', 'ParameterServerCommunicateOp',
                                                'ParameterServerSparsePullOp') for i in opt_node.inputs):
            continue
        m = _match(opt_node)
        if m is None:
            continue
        # the interior nodes must not be needed by anything outside the matched forward / backward graph
        wanted = set(id(n) for n in eval_nodes)
        if any(id(n) in wanted for n in m['interior']):
            continue
        cls = _make_op_class()
        node = cls(m, ctx=getattr(m['loss'], 'raw_ctx', None))
        loss, pred = m['loss'], m['pred']
        loss.inputs = [node]
        loss.compute = (lambda input_vals, output_val=None, stream_handle=None: input_vals[0][0])
        loss.infer_shape = (lambda input_shapes: (1,))
        pred.inputs = [node]
        pred.compute = (lambda input_vals, output_val=None, stream_handle=None: input_vals[0][1])
        pred.infer_shape = (lambda input_shapes: None)
        opt_node.inputs = [node]
        opt_node.compute = (lambda input_vals, output_val=None, stream_handle=None, new_tensors_map=None,
                            _n=node, _o=opt_node: _n.apply_update(_o))
        opt_node.wdl_fused = node
        node.opt_node = opt_node
        fused += 1
    return fused
