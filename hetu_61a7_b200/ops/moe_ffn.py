"""MoE exchange overlapped with the expert FFN: one op for dispatch -> experts -> combine (K3b/K3c/K3e).

reference chain (python/hetu/layers/moe_layer.py:60-88): layout_transform -> alltoall -> per-expert
(matmul, relu, matmul) -> concatenate -> alltoall -> reverse_layout_transform, every stage waiting for the
previous one; the two all-to-alls move 256 MiB per GPU while the tensor cores idle.

Here the exchange is cut per local expert and runs on a high-priority side stream:

    side :  barrier  disp(e0) barrier  disp(e1) barrier  zero(dY)      barrier comb(e0)    barrier comb(e1)
    main :                     wait -> h0 = relu(R0 W1) ; S0 = h0 W2    wait -> h1, S1 ...              join

``disp`` stores token rows straight into the owner GPU's receive buffer over NVLink
(csrc/moe.cu ``hetu_moe_dispatch_peer_x``), the second GEMM writes its result directly into the symmetric
send buffer (no staging copy, no concatenate), ``comb`` reads the routed rows back from the owners and
accumulates them over the chunks.  The receive buffer is never cleared: rows of unused capacity slots hold
stale finite values that no route reads; the buffer of the backward exchange is cleared under the forward
GEMMs so that such slots contribute exact zeros to the weight gradients.  The backward op mirrors the
schedule (gated dispatch of dOut, dW2 / dH / dW1 / dX GEMMs per expert, un-gated combine of dX, gate
gradients from the owners' forward outputs on the side stream).

Without a symmetric heap (CPU, one GPU, NCCL baseline) the same op runs the stages one after the other
through ``moe_dispatch`` / ``all-to-all`` / ``moe_combine``.
"""
from __future__ import annotations

import os

import torch

from .. import _lib as L
from ..graph.node import Op
from .linalg import gemm, ACT_RELU, ACT_RELU_MASK, get_matmul_precision
from .moe import moe_dispatch, moe_combine, moe_gate_grad, _ptrs
from .moe_fused import _engine, _world

_SIDE = {}


def _side_stream(device):
    """One high-priority stream per device for the exchange chunks (its CTAs are placed first when GEMM CTAs retire)."""
    key = (device.type, device.index)
    if key not in _SIDE:
        _SIDE[key] = torch.cuda.Stream(device=device, priority=-1)
    return _SIDE[key]


def _flat(ts):
    return [t.contiguous().reshape(-1) for t in ts]


def _relu_grad_(h, dh):
    """dh *= (h > 0), in place."""
    if L.kernels_on(h, dh):
        from .elementwise import B
        L.call('binary', L.c_int(B['RELU_GRAD']), L.p(h), L.p(dh), L.p(dh), L.c_sz(dh.numel()), L.c_f32(0.0),
               L.stream_ptr(dh.device))
    else:
        dh.mul_((h > 0).to(dh.dtype))
    return dh


def _gemm_into(dst, a, b, ta=False, tb=False, act=0):
    r = gemm(a, b, ta, tb, act=act, out=dst)
    if r.data_ptr() != dst.data_ptr():
        dst.copy_(r.view(dst.shape))
    return dst


class _Exchange:
    """Chunked dispatch / combine against one pair of symmetric buffers laid out [E, parts, W, C / parts, d]."""

    def __init__(self, eng, key, E, W, C, d, parts):
        self.eng, self.E, self.W, self.C, self.d, self.parts = eng, E, W, C, d, parts
        n = E * W * C * d
        self.recv_buf = eng.buffer((key, 'recv'), n * 4)
        self.send_buf = eng.buffer((key, 'send'), n * 4)
        self.recv = self.recv_buf.tensor(numel=n).view(E, W * C, d)
        self.send = self.send_buf.tensor(numel=n).view(E, W * C, d)
        self.rows = W * C // parts                  # rows of one chunk
        self.chunks = [(e, h) for e in range(E) for h in range(parts)]

    def rows_of(self, t, e, h):
        """Rows of chunk (e, h) in a per-expert [E, W*C, *] tensor (or list of per-expert tensors)."""
        return t[e][h * self.rows:(h + 1) * self.rows]

    def dispatch(self, x, idxs, locs, gates, e, h):
        k = len(idxs)
        L.call('moe_dispatch_peer_x', L.p(x), _ptrs(idxs), _ptrs(locs), _ptrs(gates or []), L.c_int(k),
               L.c_int(int(bool(gates))), self.recv_buf.ptrs, L.c_int(self.W), L.c_int(self.eng.rank),
               L.c_int(x.shape[0]), L.c_int(self.d), L.c_int(self.C), L.c_int(self.E), L.c_int(e), L.c_int(self.parts),
               L.c_int(h), L.stream_ptr(x.device))

    def combine(self, out, idxs, locs, gates, e, h, accumulate):
        k = len(idxs)
        L.call('moe_combine_peer_x', _ptrs(idxs), _ptrs(locs), _ptrs(gates or []), L.c_int(k),
               L.c_int(int(bool(gates))), self.send_buf.ptrs, L.c_int(self.W), L.c_int(self.eng.rank), L.p(out),
               L.c_int(out.shape[0]), L.c_int(self.d), L.c_int(self.C), L.c_int(self.E), L.c_int(e), L.c_int(self.parts),
               L.c_int(h), L.c_int(int(accumulate)), L.stream_ptr(out.device))


def _parts(capacity):
    """Capacity parts per expert (HETU_MOE_CHUNKS, default 1).  More chunks expose less of the first / last exchange but
    shrink the GEMMs: at 16 K rows per expert two parts cost more in tile-wave quantisation (8 K x 2 K products fill
    3.5 waves of the 74 CTA pairs) and extra barriers than they hide - measured 3.74 ms vs 3.40 ms per step on 2 GPUs."""
    want = max(1, int(os.environ.get('HETU_MOE_CHUNKS', '1')))
    while want > 1 and capacity % want:
        want -= 1
    return want


class FusedMoEFFNOp(Op):
    """(x [T, d], idx_j.., loc_j.., gate_j.., W1_e.., W2_e..) -> [T, d]: sum_j gate_j * FFN_{e_j}(x)."""

    def __init__(self, x, indices_s, location_s, gates, w1s, w2s, capacity, world, ctx=None):
        gates = list(gates) if gates is not None else []
        self.k, self.E = len(indices_s), len(w1s)
        super().__init__(FusedMoEFFNOp, [x] + list(indices_s) + list(location_s) + gates + list(w1s) + list(w2s), ctx)
        self.capacity, self.world, self.has_gate = int(capacity), int(world), len(gates) > 0
        self.hidden = None          # per-expert post-relu activations, kept for the backward op
        self.recv = None            # [E, W*C, d] expert inputs
        self.ex = None              # _Exchange (symmetric-heap path)
        self.local_rows = None      # fall-back path: expert outputs back at the source rank

    def forward_hook(self, config):
        super().forward_hook(config)
        self._config = config

    def _split(self, vals):
        k, E = self.k, self.E
        x = vals[0]
        idxs, locs = vals[1:1 + k], vals[1 + k:1 + 2 * k]
        o = 1 + 2 * k
        gates = vals[o:o + k] if self.has_gate else None
        o += k if self.has_gate else 0
        return x, idxs, locs, gates, vals[o:o + E], vals[o + E:o + 2 * E]

    def compute(self, input_vals, output_val=None, stream_handle=None):
        x, idxs, locs, gates, w1s, w2s = self._split(input_vals)
        E, W, C = self.E, self.world, self.capacity
        T, d = x.shape
        if x.device.type == 'meta':
            return torch.empty((T, d), device='meta')
        eng, comm = _engine(self), _world(self)
        idxs, locs = _flat(idxs), _flat(locs)
        gates = _flat(gates) if gates else None
        x = x.contiguous()
        if eng is not None and comm is not None and x.is_cuda and d % 4 == 0:
            return self._overlapped(eng, x, idxs, locs, gates, w1s, w2s)
        # one stage after the other
        local = moe_dispatch(x, idxs, locs, C, E * W)                          # [W*E*C, d], destination-rank major
        if comm is not None:
            local = comm.alltoall(local.contiguous())                         # [W(src), E, C, d]
        recv = local.view(W, E, C, d).permute(1, 0, 2, 3).contiguous().view(E, W * C, d)
        self.recv, self.hidden, self.ex = recv, [], None
        ys = torch.empty((E, W * C, d), dtype=x.dtype, device=x.device)
        for e in range(E):
            h = gemm(recv[e], w1s[e], act=ACT_RELU)
            self.hidden.append(h)
            _gemm_into(ys[e], h, w2s[e])
        rows = ys.view(E, W, C, d).permute(1, 0, 2, 3).contiguous().view(W * E * C, d)
        if comm is not None:
            rows = comm.alltoall(rows)
        self.local_rows = rows
        return moe_combine(rows, idxs, locs, C, gates=gates)

    def _overlapped(self, eng, x, idxs, locs, gates, w1s, w2s):
        E, W, C = self.E, self.world, self.capacity
        T, d = x.shape
        if self.ex is None:
            parts = _parts(C)
            self.ex = _Exchange(eng, ('moe_ffn', self.id), E, W, C, d, parts)
            self.bwd_ex = _Exchange(eng, ('moe_ffn_bwd', self.id), E, W, C, d, parts)
        ex = self.ex
        self.recv = ex.recv
        main, side = torch.cuda.current_stream(x.device), _side_stream(x.device)
        out = torch.empty((T, d), dtype=torch.float32, device=x.device)
        hid = w1s[0].shape[1]
        if self.hidden is None or len(self.hidden) != E or self.hidden[0].shape != (W * C, hid):
            self.hidden = [torch.empty((W * C, hid), dtype=torch.float32, device=x.device) for _ in range(E)]
        n = len(ex.chunks)
        arrived = [torch.cuda.Event() for _ in range(n)]
        produced = [torch.cuda.Event() for _ in range(n)]
        fork, done = torch.cuda.Event(), torch.cuda.Event()
        fork.record(main)
        with torch.cuda.stream(side):
            side.wait_event(fork)
            eng.barrier()                          # every rank is past the last reader of its receive buffer
            for c, (e, h) in enumerate(ex.chunks):
                ex.dispatch(x, idxs, locs, None, e, h)
                eng.barrier()                      # this chunk of every peer has landed
                arrived[c].record(side)
            self.bwd_ex.recv.zero_()               # under the GEMMs: unused slots of dY must be exact zeros
        for c, (e, h) in enumerate(ex.chunks):
            main.wait_event(arrived[c])
            hc = ex.rows_of(self.hidden, e, h)
            _gemm_into(hc, ex.rows_of(ex.recv, e, h), w1s[e], act=ACT_RELU)
            _gemm_into(ex.rows_of(ex.send, e, h), hc, w2s[e])
            produced[c].record(main)
        with torch.cuda.stream(side):
            for c, (e, h) in enumerate(ex.chunks):
                side.wait_event(produced[c])
                eng.barrier()                      # every rank's outputs of this chunk are in place
                ex.combine(out, idxs, locs, gates, e, h, accumulate=c > 0)
            done.record(side)
        main.wait_event(done)
        return out

    def gradient(self, g):
        grad = FusedMoEFFNGradOp(self, g, ctx=self.raw_ctx)
        k, E = self.k, self.E
        outs = [_MoEPick(grad, self.inputs[0], 'dx', ctx=self.raw_ctx)] + [None] * (2 * k)
        o = 1 + 2 * k
        if self.has_gate:
            outs += [_MoEPick(grad, self.inputs[o + j], ('dgate', j), ctx=self.raw_ctx) for j in range(k)]
            o += k
        outs += [_MoEPick(grad, self.inputs[o + e], ('dw1', e), ctx=self.raw_ctx) for e in range(E)]
        outs += [_MoEPick(grad, self.inputs[o + E + e], ('dw2', e), ctx=self.raw_ctx) for e in range(E)]
        return outs

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[0])


class FusedMoEFFNGradOp(Op):
    """All gradients of FusedMoEFFNOp in one schedule; ``_MoEPick`` nodes hand the pieces to the graph."""

    def __init__(self, fwd, grad, ctx=None):
        super().__init__(FusedMoEFFNGradOp, [grad, fwd] + list(fwd.inputs), ctx)
        self.forward_node = fwd
        self.results = None
        self.wanted = set()         # filled by the _MoEPick nodes that are part of the executed graph

    def forward_hook(self, config):
        super().forward_hook(config)
        self._config = config

    def compute(self, input_vals, output_val=None, stream_handle=None):
        fwd = self.forward_node
        g = input_vals[0]
        need_dx = 'dx' in self.wanted or not self.wanted
        x, idxs, locs, gates, w1s, w2s = fwd._split(input_vals[2:])
        if g.device.type == 'meta':
            return torch.empty_like(x if need_dx else w1s[0])
        E, W, C, k = fwd.E, fwd.world, fwd.capacity, fwd.k
        T, d = x.shape
        idxs, locs = _flat(idxs), _flat(locs)
        gates = _flat(gates) if gates else None
        g = g.contiguous()
        res = {}
        if fwd.ex is not None and _engine(self) is not None:
            self._overlapped(_engine(self), fwd, g, idxs, locs, gates, w1s, w2s, res, need_dx)
        else:
            comm = _world(self)
            local = moe_dispatch(g, idxs, locs, C, E * W, gates=gates)
            if comm is not None:
                local = comm.alltoall(local.contiguous())
            dy = local.view(W, E, C, d).permute(1, 0, 2, 3).contiguous().view(E, W * C, d)
            dxs = torch.empty((E, W * C, d), dtype=g.dtype, device=g.device) if need_dx else None
            for e in range(E):
                self._expert_grads(fwd, e, dy[e], w1s[e], w2s[e], res, dxs[e] if need_dx else None)
            if need_dx:
                rows = dxs.view(E, W, C, d).permute(1, 0, 2, 3).contiguous().view(W * E * C, d)
                if comm is not None:
                    rows = comm.alltoall(rows)
                res['dx'] = moe_combine(rows, idxs, locs, C, gates=None)
            if gates:
                for j in range(k):
                    res[('dgate', j)] = moe_gate_grad(g, fwd.local_rows, idxs[j], locs[j], C)
        self.results = res
        return res['dx'] if need_dx else res[('dw1', 0)]

    @staticmethod
    def _masked_dh(h, dy, w2):
        """dH = (dY W2^T) * (h > 0); over h when the tensor-core epilogue can do it."""
        m, n, kk = dy.shape[0], w2.shape[0], dy.shape[1]
        if (L.kernels_on(h, dy, w2) and get_matmul_precision() == 'tf32' and m >= 512 and n >= 256 and kk >= 32
                and n % 4 == 0 and kk % 4 == 0 and float(m) * n * kk > float(1 << 25) and h.is_contiguous()):
            return gemm(dy, w2, False, True, act=ACT_RELU_MASK, out=h)
        return _relu_grad_(h, gemm(dy, w2, False, True))

    @classmethod
    def _expert_grads(cls, fwd, e, dy, w1, w2, res, dx_out):
        h = fwd.hidden[e]
        res[('dw2', e)] = gemm(h, dy, True, False)               # h^T dY
        dh = cls._masked_dh(h, dy, w2)
        res[('dw1', e)] = gemm(fwd.recv[e], dh, True, False)     # R^T dH
        if dx_out is not None:                                   # the layer input may need no gradient (first layer)
            _gemm_into(dx_out, dh, w1, False, True)              # dH W1^T

    def _overlapped(self, eng, fwd, g, idxs, locs, gates, w1s, w2s, res, need_dx):
        E, k = fwd.E, fwd.k
        ex = fwd.bwd_ex
        T, d = g.shape
        main, side = torch.cuda.current_stream(g.device), _side_stream(g.device)
        dx = torch.empty((T, d), dtype=torch.float32, device=g.device) if need_dx else None
        dgate = [torch.empty((T,), dtype=torch.float32, device=g.device) for _ in range(k)] if gates else []
        n = len(ex.chunks)
        arrived = [torch.cuda.Event() for _ in range(n)]
        produced = [torch.cuda.Event() for _ in range(n)]
        fork, done = torch.cuda.Event(), torch.cuda.Event()
        fork.record(main)
        with torch.cuda.stream(side):
            side.wait_event(fork)
            for c, (e, h) in enumerate(ex.chunks):     # dY buffer was cleared (and a barrier passed) in the forward op
                ex.dispatch(g, idxs, locs, gates, e, h)
                eng.barrier()
                arrived[c].record(side)
            for j in range(k if gates else 0):     # d gate_j[t] = <dOut[t], expert output row of route j>, read at its owner
                L.call('moe_gate_grad_peer_x', L.p(g), L.p(idxs[j]), L.p(locs[j]), fwd.ex.send_buf.ptrs, L.c_int(fwd.world),
                       L.c_int(eng.rank), L.p(dgate[j]), L.c_int(T), L.c_int(d), L.c_int(fwd.capacity), L.c_int(E),
                       L.c_int(ex.parts), L.stream_ptr(g.device))
        for c, (e, h) in enumerate(ex.chunks):
            main.wait_event(arrived[c])
            hc, dyc = ex.rows_of(fwd.hidden, e, h), ex.rows_of(ex.recv, e, h)
            if h == 0:
                res[('dw2', e)] = gemm(hc, dyc, True, False)                          # h^T dY, summed over the expert's chunks
            else:
                gemm(hc, dyc, True, False, c_in=res[('dw2', e)], beta=1.0, out=res[('dw2', e)])
            dh = self._masked_dh(hc, dyc, w2s[e])
            if dh.data_ptr() != hc.data_ptr():
                hc.copy_(dh)                                                          # small shapes: no in-place epilogue
            if need_dx:
                _gemm_into(ex.rows_of(ex.send, e, h), hc, w1s[e], False, True)        # dH W1^T
            produced[c].record(main)
            if h == ex.parts - 1:                                                     # dH of the whole expert is in fwd.hidden[e] now
                res[('dw1', e)] = gemm(fwd.recv[e], fwd.hidden[e], True, False)       # R^T dH
        with torch.cuda.stream(side):
            for c, (e, h) in enumerate(ex.chunks if need_dx else []):
                side.wait_event(produced[c])
                eng.barrier()
                ex.combine(dx, idxs, locs, None, e, h, accumulate=c > 0)
            done.record(side)
        main.wait_event(done)
        if need_dx:
            res['dx'] = dx
        for j in range(len(dgate)):
            res[('dgate', j)] = dgate[j]

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[2]) if ('dx' in self.wanted or not self.wanted) else tuple(input_shapes[-2 * self.forward_node.E])


class _MoEPick(Op):
    def __init__(self, grad_node, like, which, ctx=None):
        super().__init__(_MoEPick, [grad_node, like], ctx)
        self.which = which
        self.op_type = 'MoEFFNGradPick'

    def forward_hook(self, config):
        super().forward_hook(config)
        self.inputs[0].wanted.add(self.which)

    def compute(self, input_vals, output_val=None, stream_handle=None):
        like = input_vals[1]
        if like.device.type == 'meta':
            return torch.empty_like(like)
        return self.inputs[0].results[self.which].view(like.shape)

    def infer_shape(self, input_shapes):
        return tuple(input_shapes[1])


def fused_moe_ffn_op(x, indices_s, location_s, gates, w1s, w2s, capacity, world, ctx=None):
    return FusedMoEFFNOp(x, indices_s, location_s, gates, w1s, w2s, capacity, world, ctx=ctx)
