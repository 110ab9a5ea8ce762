"""GBDT / random-forest training on the GPU (``GBDT(device='gpu')``).

The CPU core (csrc_ml/ml.cc) grows a tree node by node; on a B200 the unit of work is a tree
LEVEL: ``hetu_gbdt_hist`` (csrc/gbdt.cu) builds the histograms of all open nodes of the level in
one launch, split search is a handful of vectorised scans over the ``[nodes, bins]`` histogram
tensor on the device, and ``hetu_gbdt_route`` moves every instance to its child.  Cut points and
bins come from the same native routines as the CPU trainer, and the finished trees are handed to
the native model (text format of ``hml_gbdt_save``), so prediction / save / load are shared.

Supported here: level-wise growth with scalar leaves (binary, regression, one tree per class for
multi-class).  Leaf-wise growth and vector-leaf trees stay on the CPU trainer."""
from __future__ import annotations

import ctypes
import os
import tempfile

import numpy as np
import torch

from .. import _lib as L
from . import _core as C


def supported(args) -> bool:
    multi = (not args['is_regression']) and args['num_label'] > 2
    return (not args['leafwise']) and (not multi or bool(args['multi_tree']))


def _cuts_and_bins(X, dim, num_split):
    ip, idx, val = X.ptrs()
    n = X.get_num_rows()
    off = np.zeros(dim + 1, dtype=np.int64)
    cap = int(dim) * (int(num_split) + 2) + 8
    cut_vals = np.zeros(cap, dtype=np.float32)
    lib = C.lib()
    lib.hml_gbdt_cuts.restype = ctypes.c_int64
    got = lib.hml_gbdt_cuts(ip, idx, val, ctypes.c_int64(n), ctypes.c_int64(dim), int(num_split), off.ctypes.data_as(C.I64P),
                            cut_vals.ctypes.data_as(C.F32P), ctypes.c_int64(cap))
    if got < 0:
        raise RuntimeError('hml_gbdt_cuts failed (%d): %s' % (got, lib.hml_last_error().decode()))
    cut_vals = cut_vals[:got]
    bins = np.zeros(X.indices.size, dtype=np.uint8)
    C.check(lib.hml_gbdt_bins(ip, idx, val, ctypes.c_int64(n), ctypes.c_int64(dim), off.ctypes.data_as(C.I64P),
                              cut_vals.ctypes.data_as(C.F32P), bins.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8))), 'hml_gbdt_bins')
    return off, cut_vals, bins


class _Level(object):
    """Device-side constants of one data set."""

    def __init__(self, X, dim, num_split, dev):
        off, cut_vals, bins = _cuts_and_bins(X, dim, num_split)
        ncut = np.diff(off)
        self.dim = dim
        self.cut_off, self.cut_vals = off, cut_vals
        bin_off = off + np.arange(dim + 1)                       # feature f owns ncut_f + 1 bins
        self.total_bins = int(bin_off[-1])
        zero_bin = np.array([np.searchsorted(cut_vals[off[f]:off[f + 1]], 0.0, side='left') for f in range(dim)], dtype=np.int32)
        feat_of_bin = np.repeat(np.arange(dim), ncut + 1)
        last_bin = np.zeros(self.total_bins, dtype=bool)
        last_bin[bin_off[1:] - 1] = True
        t = lambda a, dt: torch.as_tensor(np.ascontiguousarray(a), dtype=dt, device=dev)     # noqa: E731
        self.indptr, self.indices = t(X.indptr, torch.int64), t(X.indices, torch.int32)
        self.bins = t(bins, torch.uint8)
        self.bin_off, self.zero_bin = t(bin_off, torch.int64), t(zero_bin, torch.int32)
        self.feat_of_bin, self.last_bin = t(feat_of_bin, torch.int64), t(last_bin, torch.bool)
        self.zero_pos = self.bin_off[:-1] + self.zero_bin.long()  # histogram column of each feature's zero bin
        self.first_bin = self.bin_off[:-1]
        self.end_bin = self.bin_off[1:] - 1
        self.usable_feat = t(ncut > 0, torch.bool)
        self.n = X.get_num_rows()
        self.dev = dev


def _soft(G, alpha):
    return torch.sign(G) * torch.clamp(G.abs() - alpha, min=0.0) if alpha > 0 else G


def _grow_tree(lv: _Level, g, h, p, feat_on, leaf_scale):
    """One level-wise tree.  Returns (nodes list [feat, thr, left, right, leaf], leaf values, node_of)."""
    dev, n, TB = lv.dev, lv.n, lv.total_bins
    lam, alpha = float(p['reg_lambda']), float(p['reg_alpha'])
    node_of = torch.zeros(n, dtype=torch.int32, device=dev)
    feat_on_u8 = feat_on.to(torch.uint8)
    nodes = [[-1, 0.0, -1, -1, -1]]
    leaf_vals = []
    open_nodes = [0]
    bin_feat_on = feat_on[lv.feat_of_bin]
    for depth in range(int(p['max_depth']) + 1):
        if not open_nodes:
            break
        A = len(open_nodes)
        slot_of_node = torch.full((len(nodes),), -1, dtype=torch.int32, device=dev)
        slot_of_node[torch.as_tensor(open_nodes, device=dev)] = torch.arange(A, dtype=torch.int32, device=dev)
        slots = slot_of_node[node_of.long()].long()
        live = slots >= 0
        tot = torch.zeros((A, 3), dtype=torch.float32, device=dev)
        tot.index_add_(0, slots[live], torch.stack([g[live], h[live], (h[live] != 0).float()], dim=1))
        cnt_all = torch.zeros(A, dtype=torch.float32, device=dev)
        cnt_all.index_add_(0, slots[live], torch.ones_like(g[live]))
        can_split = depth < int(p['max_depth'])
        best_gain = torch.full((A,), -1.0, device=dev)
        best_bin = torch.zeros(A, dtype=torch.int64, device=dev)
        if can_split:
            hist = torch.zeros((A, TB, 2), dtype=torch.float32, device=dev)
            if dev.type == 'cuda':
                L.call('gbdt_hist', L.p(lv.indptr), L.p(lv.indices), L.p(lv.bins), L.p(node_of), L.p(slot_of_node), L.p(g), L.p(h),
                       L.p(lv.bin_off), L.p(feat_on_u8), L.p(hist), L.c_i64(TB), L.c_i64(n), L.stream_ptr(dev))
            else:
                _hist_reference(lv, node_of, slot_of_node, g, h, feat_on, hist)
            # zero-valued rows: node total minus the visited mass of the feature goes to the feature's zero bin
            cs = hist.cumsum(dim=1)
            before = torch.where((lv.first_bin > 0).view(1, -1, 1), cs[:, (lv.first_bin - 1).clamp(min=0)], torch.zeros_like(cs[:, :1]))
            fsum = cs[:, lv.end_bin] - before                                            # [A, dim, 2]
            resid = (tot[:, None, :2] - fsum) * feat_on.view(1, -1, 1)
            hist.index_add_(1, lv.zero_pos, resid)
            cs = hist.cumsum(dim=1)
            before = torch.where((lv.first_bin > 0).view(1, -1, 1), cs[:, (lv.first_bin - 1).clamp(min=0)], torch.zeros_like(cs[:, :1]))
            left = cs - before[:, lv.feat_of_bin]                                        # [A, TB, 2] prefix inside the feature
            GL, HL = left[..., 0], left[..., 1]
            GR, HR = tot[:, None, 0] - GL, tot[:, None, 1] - HL
            parent = _soft(tot[:, 0], alpha) ** 2 / (tot[:, 1] + lam + 1e-12)
            gain = 0.5 * (_soft(GL, alpha) ** 2 / (HL + lam + 1e-12) + _soft(GR, alpha) ** 2 / (HR + lam + 1e-12) - parent[:, None])
            mcw = float(p['min_child_weight'])
            bad = lv.last_bin.view(1, -1) | ~bin_feat_on.view(1, -1) | (HL <= 0) | (HR <= 0) | (HL < mcw) | (HR < mcw)
            gain = torch.where(bad, torch.full_like(gain, -1.0), gain)
            best_gain, best_bin = gain.max(dim=1)
        host = torch.cat([tot, cnt_all[:, None], best_gain[:, None], best_bin[:, None].float()], dim=1).cpu().numpy()
        split_feat = np.full(len(nodes), -1, dtype=np.int32)
        split_bin = np.zeros(len(nodes), dtype=np.int32)
        left_child = np.zeros(len(nodes), dtype=np.int32)
        next_open = []
        bin_off_h = None
        for a, nd in enumerate(open_nodes):
            G, H, cnt, _, gain_a, bb = host[a]                     # cnt = sampled instances of the node, as on the CPU
            do = (can_split and gain_a > float(p['min_split_gain']) and gain_a > 0 and cnt >= float(p['min_node_ins']) and cnt >= 2
                  and len(nodes) + 2 <= int(p['max_node_num']))
            if do:
                if bin_off_h is None:
                    bin_off_h = lv.bin_off.cpu().numpy()
                f = int(np.searchsorted(bin_off_h, int(bb), side='right') - 1)
                b = int(bb) - int(bin_off_h[f])
                li = len(nodes)
                nodes.append([-1, 0.0, -1, -1, -1])
                nodes.append([-1, 0.0, -1, -1, -1])
                nodes[nd][0], nodes[nd][1], nodes[nd][2], nodes[nd][3] = f, float(lv.cut_vals[lv.cut_off[f] + b]), li, li + 1
                split_feat[nd], split_bin[nd], left_child[nd] = f, b, li
                next_open += [li, li + 1]
            else:
                t = np.sign(G) * max(abs(G) - alpha, 0.0) if alpha > 0 else G
                w = -t / (H + lam + 1e-12)
                if float(p['max_leaf_weight']) > 0:
                    w = float(np.clip(w, -p['max_leaf_weight'], p['max_leaf_weight']))
                nodes[nd][4] = len(leaf_vals)
                leaf_vals.append(float(w * leaf_scale))
        if next_open:
            pad = len(nodes) - len(split_feat)
            sf = torch.as_tensor(np.concatenate([split_feat, np.full(pad, -1, np.int32)]), device=dev)
            sb = torch.as_tensor(np.concatenate([split_bin, np.zeros(pad, np.int32)]), device=dev)
            lc = torch.as_tensor(np.concatenate([left_child, np.zeros(pad, np.int32)]), device=dev)
            if dev.type == 'cuda':
                L.call('gbdt_route', L.p(lv.indptr), L.p(lv.indices), L.p(lv.bins), L.p(node_of), L.p(sf), L.p(sb), L.p(lv.zero_bin),
                       L.p(lc), L.c_i64(n), L.stream_ptr(dev))
            else:
                _route_reference(lv, node_of, sf, sb, lc)
        open_nodes = next_open
    leaf_of_node = torch.as_tensor([nd[4] for nd in nodes], dtype=torch.int64, device=dev)
    return nodes, leaf_vals, leaf_of_node[node_of.long()]


def _hist_reference(lv, node_of, slot_of_node, g, h, feat_on, hist):
    """What hetu_gbdt_hist computes, in torch (host oracle for the kernel and for the level logic)."""
    row = torch.repeat_interleave(torch.arange(lv.n), lv.indptr[1:] - lv.indptr[:-1])
    f = lv.indices.long()
    slot = slot_of_node[node_of[row].long()].long()
    ok = (slot >= 0) & feat_on[f] & ~((g[row] == 0) & (h[row] == 0))
    pos = slot[ok] * lv.total_bins + lv.bin_off[f[ok]] + lv.bins[ok].long()
    flat = hist.view(-1, 2)
    flat.index_add_(0, pos, torch.stack([g[row][ok], h[row][ok]], dim=1))


def _route_reference(lv, node_of, split_feat, split_bin, left_child):
    indptr, indices, bins = lv.indptr.numpy(), lv.indices.numpy(), lv.bins.numpy()
    sf, sb, lc, zb = split_feat.numpy(), split_bin.numpy(), left_child.numpy(), lv.zero_bin.numpy()
    no = node_of.numpy()
    for i in range(lv.n):
        f = sf[no[i]]
        if f < 0:
            continue
        row = indices[indptr[i]:indptr[i + 1]]
        k = np.searchsorted(row, f)
        b = bins[indptr[i] + k] if k < row.size and row[k] == f else zb[f]
        no[i] = lc[no[i]] if b <= sb[no[i]] else lc[no[i]] + 1


def fit(model, train, dim, dev=None):
    """Train on the GPU and load the result into ``model``'s native handle (``dev='cpu'`` runs the
    same level logic with torch stand-ins for the two kernels - the host oracle used by the tests)."""
    p = model._args
    dev = torch.device(dev or 'cuda')
    X = train.X
    K = p['num_label'] if (not p['is_regression'] and p['num_label'] > 2) else 1
    y_np = train.y if (p['is_regression'] or K > 1) else train.labels01()
    lv = _Level(X, dim, int(p['num_split']), dev)
    n = lv.n
    y = torch.as_tensor(y_np, dtype=torch.float32, device=dev)
    score = torch.zeros((n, K), dtype=torch.float32, device=dev)
    gen = torch.Generator(device='cpu').manual_seed(int(p.get('seed', 42)))
    majority = bool(p.get('majority_voting', False))
    lr = 1.0 if majority else float(p['learning_rate'])
    trees, tree_class, curve = [], [], []
    for r in range(int(p['num_round'])):
        s = torch.zeros_like(score) if majority else score
        if K == 1:
            if p['is_regression']:
                g_all, h_all = (s[:, 0] - y)[:, None], torch.ones((n, 1), device=dev)
            else:
                pr = torch.sigmoid(s[:, 0])
                g_all, h_all = (pr - y)[:, None], torch.clamp(pr * (1 - pr), min=1e-16)[:, None]
        else:
            pr = torch.softmax(s, dim=1)
            g_all = pr - torch.nn.functional.one_hot(y.long(), K).float()
            h_all = torch.clamp(pr * (1 - pr), min=1e-16)
        keep = torch.ones(n, device=dev)
        if p['ins_sp_ratio'] < 1:
            keep = (torch.rand(n, generator=gen) < p['ins_sp_ratio']).to(dev).float()
        feat_on = lv.usable_feat.clone()
        if p['feat_sp_ratio'] < 1:
            feat_on &= (torch.rand(dim, generator=gen) < p['feat_sp_ratio']).to(dev)
        for k in range(K):
            g = (g_all[:, k] * keep).contiguous()
            h = (h_all[:, k] * keep).contiguous()
            nodes, leaf_vals, leaf_idx = _grow_tree(lv, g, h, p, feat_on, lr)
            lvals = torch.as_tensor(leaf_vals, dtype=torch.float32, device=dev)
            score[:, k] += lvals[leaf_idx]
            trees.append((nodes, leaf_vals))
            tree_class.append(k if K > 1 else -1)
        sc = score / (r + 1) if majority else score
        if K == 1 and p['is_regression']:
            loss = torch.mean((sc[:, 0] - y) ** 2)
        elif K == 1:
            loss = torch.nn.functional.binary_cross_entropy_with_logits(sc[:, 0], y)
        else:
            loss = torch.nn.functional.cross_entropy(sc, y.long())
        curve.append(float(loss))
    _load_into_native(model, trees, tree_class, dim, K, curve)
    return curve


def _load_into_native(model, trees, tree_class, dim, K, curve):
    from .ensemble import _ORDER
    p = dict(model._args)
    if K > 1:
        p['multi_tree'] = True
    fd, path = tempfile.mkstemp(suffix='.gbdt')
    try:
        with os.fdopen(fd, 'w') as f:
            f.write('hetuml-gbdt 1\n%d %d %d\n' % (dim, K, len(trees)))
            vals = []
            for k in _ORDER:
                v = p[k]
                vals.append('%d' % int(v) if isinstance(v, (bool, int, np.integer)) else '%.17g' % float(v))
            f.write(' '.join(vals) + '\n')
            for (nodes, leaf_vals), tc in zip(trees, tree_class):
                f.write('%d 1 %d %d\n' % (tc, len(nodes), len(leaf_vals)))
                for nd in nodes:
                    f.write('%d %.9g %d %d %d\n' % (nd[0], nd[1], nd[2], nd[3], nd[4]))
                f.write(' '.join('%.9g' % v for v in leaf_vals) + '\n')
        model._new_handle()
        C.check(C.lib().hml_gbdt_load(ctypes.c_void_p(model._handle), path.encode()), 'GBDT load')
    finally:
        os.unlink(path)
    model._gpu_curve = np.asarray(curve)
