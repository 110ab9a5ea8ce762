#!/usr/bin/env python
"""Layout search for transformer training on one NVSwitch node: enumerate (pipeline, tensor, data)
parallel degrees and micro-batch counts, estimate step time and per-GPU memory with an analytic
model, print the ranking and the command line for ``examples/nlp/bert/train_hetu_bert_hybrid.py``.

reference: tools/Galvatron/README.md is a two-line placeholder in the reference snapshot (no code);
this is a small working counterpart for the layouts this framework can run (PP x TP(FFN) x DP).

Cost model (per training step, one replica = pp * tp GPUs):
  compute      6 * params_per_stage * tokens / (tp * peak * mfu)               fwd + bwd GEMM FLOPs
  tp comm      4 all-reduces of [tokens, hidden] per layer (fwd + bwd), 2(tp-1)/tp * bytes / nvlink_bw
  pp bubble    (pp - 1) / micro_batches of the per-stage time; p2p activations [tokens, hidden] per boundary
  dp comm      2(dp-1)/dp * grad_bytes / nvlink_bw, overlapped with backward up to `overlap`
  memory       weights + grads + Adam slots (16 B / param) / (tp * pp) + activations of the in-flight micro-batches
Peaks come from MEASURED_PEAKS.json when present (driver-written), else B200 nominal values.

    python tools/Galvatron/search.py --gpus 8 --model bert-large --global-batch 256 --seq-len 128
"""
import argparse
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
MODELS = {'bert-base': dict(hidden=768, layers=12, heads=12, ffn=3072, vocab=30522),
          'bert-large': dict(hidden=1024, layers=24, heads=16, ffn=4096, vocab=30522)}


def peaks():
    p = dict(tflops=1400.0, hbm_gbs=6500.0, nvlink_gbs=750.0)      # sustained bf16 cuBLAS, HBM copy, NVLink per direction (achievable)
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        try:
            m = json.load(open(path))
            p['tflops'] = float(m.get('bf16_tflops_sustained', m.get('bf16_tflops', p['tflops'])))
            p['hbm_gbs'] = float(m.get('hbm_gbs', p['hbm_gbs']))
        except (ValueError, OSError):
            pass
    return p


def layer_params(m):
    return 4 * m['hidden'] * m['hidden'] + 2 * m['hidden'] * m['ffn'] + 9 * m['hidden'] + m['ffn']


def estimate(m, gpus, pp, tp, dp, micro, global_batch, seq, pk, tf32=True, mfu=0.45, overlap=0.7, hbm_gb=180.0):
    if m['layers'] % pp or global_batch % (dp * micro):
        return None
    mb = global_batch // (dp * micro)                                 # sequences per micro-batch per replica
    tokens = mb * seq
    peak = pk['tflops'] * 1e12 * (0.5 if tf32 else 1.0) * mfu         # tf32 tensor-core rate is half the bf16 rate
    lp = layer_params(m)
    stage_layers = m['layers'] // pp
    embed = m['vocab'] * m['hidden']
    stage_params = stage_layers * lp + (embed if pp == 1 else embed // 2)
    attn_flops = 4 * tokens * seq * m['hidden']                       # QK^T and PV per layer, fwd
    t_comp = (6 * stage_layers * lp * tokens / tp + 3 * stage_layers * attn_flops) / peak
    bw = pk['nvlink_gbs'] * 1e9
    act_bytes = tokens * m['hidden'] * 4
    t_tp = 0.0 if tp == 1 else stage_layers * 4 * (2 * (tp - 1) / tp) * act_bytes / bw
    t_stage = t_comp + t_tp
    t_pipe = t_stage * (micro + pp - 1)                               # GPipe / 1F1B: same bubble fraction
    t_p2p = 0.0 if pp == 1 else 2 * micro * act_bytes / bw
    grad_bytes = stage_params / tp * 4
    t_dp = 0.0 if dp == 1 else (2 * (dp - 1) / dp) * grad_bytes / bw * (1 - overlap)
    t_opt = 3 * stage_params / tp * 16 / (pk['hbm_gbs'] * 1e9)         # Adam: read/write weights + 2 slots + grad
    step = t_pipe + t_p2p + t_dp + t_opt
    act_per_layer = tokens * (34 * m['hidden'] + 5 * m['heads'] * seq) * 4 / max(tp, 1) / 4   # fp32 activations kept for backward
    in_flight = min(micro, pp)
    mem = (stage_params / tp * 16 + stage_layers * act_per_layer * in_flight) / 1e9
    return dict(pp=pp, tp=tp, dp=dp, micro=micro, step_ms=step * 1e3, tokens_per_s=global_batch * seq / step,
                mem_gb=mem, fits=mem < hbm_gb * 0.9, bubble=(pp - 1) / (micro + pp - 1),
                parts=dict(compute=t_comp * micro * 1e3, tp=t_tp * micro * 1e3, dp=t_dp * 1e3, p2p=t_p2p * 1e3, opt=t_opt * 1e3))


def search(model, gpus, global_batch, seq, tf32=True, micro_choices=(1, 2, 4, 8, 16)):
    m, pk = MODELS[model], peaks()
    out = []
    for pp in [d for d in range(1, gpus + 1) if gpus % d == 0]:
        for tp in [d for d in range(1, gpus // pp + 1) if (gpus // pp) % d == 0]:
            dp = gpus // (pp * tp)
            if m['heads'] % tp or m['ffn'] % tp:
                continue
            for micro in micro_choices:
                if pp == 1 and micro != 1:
                    continue
                e = estimate(m, gpus, pp, tp, dp, micro, global_batch, seq, pk, tf32)
                if e is not None:
                    out.append(e)
    return sorted(out, key=lambda e: (not e['fits'], e['step_ms']))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=8)
    ap.add_argument('--model', default='bert-large', choices=sorted(MODELS))
    ap.add_argument('--global-batch', type=int, default=256)
    ap.add_argument('--seq-len', type=int, default=128)
    ap.add_argument('--top', type=int, default=8)
    ap.add_argument('--bf16', action='store_true', help='model bf16 tensor-core rate instead of tf32')
    args = ap.parse_args()
    ranked = search(args.model, args.gpus, args.global_batch, args.seq_len, tf32=not args.bf16)
    print('%-4s %-4s %-4s %-6s %10s %14s %9s %7s  %s' % ('pp', 'tp', 'dp', 'micro', 'step ms', 'tokens/s', 'mem GB', 'bubble', 'compute / tp / dp / p2p / opt (ms)'))
    for e in ranked[:args.top]:
        p = e['parts']
        print('%-4d %-4d %-4d %-6d %10.2f %14.0f %9.1f %7.2f  %.2f / %.2f / %.2f / %.2f / %.2f%s' % (
            e['pp'], e['tp'], e['dp'], e['micro'], e['step_ms'], e['tokens_per_s'], e['mem_gb'], e['bubble'], p['compute'], p['tp'],
            p['dp'], p['p2p'], p['opt'], '' if e['fits'] else '   (does not fit)'))
    if ranked:
        b = ranked[0]
        per = args.global_batch // (b['dp'] * b['micro'])
        print('\nbest: heturun -w %d python examples/nlp/bert/train_hetu_bert_hybrid.py --size %s --pp %d --tp %d --dp %d '
              '--micro-batches %d --batch-size %d --seq-len %d' % (args.gpus, args.model.split('-')[1], b['pp'], b['tp'], b['dp'],
                                                                  b['micro'], per, args.seq_len))


if __name__ == '__main__':
    main()
