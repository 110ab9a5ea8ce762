#!/usr/bin/env python
"""Prepare a parallel corpus for the Transformer example (reference examples/nlp/prepare_data.py:
IWSLT'16 xml/tag clean-up, train / eval / test split, joint sentencepiece BPE model, segmented files).

    python prepare_data.py --src raw.de --tgt raw.en --out iwslt2016 --vocab_size 8000

Uses ``sentencepiece`` for the joint BPE model when it can train one (enough text), otherwise falls
back to a whitespace vocabulary, so the pipeline also works on toy corpora."""
import argparse
import collections
import os
import re


def clean(lines):
    out = []
    for line in lines:
        line = line.strip()
        if not line or line.startswith('<'):                  # IWSLT meta lines: <url>, <talkid>, ...
            continue
        out.append(re.sub(r'<[^>]+>', '', line).strip())
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--src', required=True)
    ap.add_argument('--tgt', required=True)
    ap.add_argument('--out', default='iwslt2016')
    ap.add_argument('--vocab_size', type=int, default=32000)
    ap.add_argument('--eval_fraction', type=float, default=0.02)
    args = ap.parse_args()
    src = clean(open(args.src, encoding='utf-8').read().splitlines())
    tgt = clean(open(args.tgt, encoding='utf-8').read().splitlines())
    n = min(len(src), len(tgt))
    src, tgt = src[:n], tgt[:n]
    n_eval = max(1, int(n * args.eval_fraction))
    seg_dir = os.path.join(args.out, 'segmented')
    os.makedirs(seg_dir, exist_ok=True)
    joint = os.path.join(args.out, 'train.joint')
    with open(joint, 'w', encoding='utf-8') as f:
        f.write('\n'.join(src[n_eval:] + tgt[n_eval:]) + '\n')
    pieces = None
    try:
        import sentencepiece as spm
        prefix = os.path.join(seg_dir, 'bpe')
        spm.SentencePieceTrainer.Train('--input=%s --pad_id=0 --unk_id=1 --bos_id=2 --eos_id=3 --model_prefix=%s '
                                       '--vocab_size=%d --model_type=bpe --hard_vocab_limit=false' % (joint, prefix, args.vocab_size))
        sp = spm.SentencePieceProcessor()
        sp.Load(prefix + '.model')
        pieces = lambda s: ' '.join(sp.EncodeAsPieces(s))     # noqa: E731
    except Exception as e:                                     # toy corpus / no sentencepiece: whitespace tokens
        print('sentencepiece unavailable or corpus too small (%s): using a whitespace vocabulary' % e)
        cnt = collections.Counter(w for s in src + tgt for w in s.split())
        vocab = ['<pad>', '<unk>', '<s>', '</s>'] + [w for w, _ in cnt.most_common(args.vocab_size - 4)]
        with open(os.path.join(seg_dir, 'bpe.vocab'), 'w', encoding='utf-8') as f:
            f.write('\n'.join(vocab) + '\n')
        pieces = lambda s: s                                   # noqa: E731
    for name, a, b in (('train', src[n_eval:], tgt[n_eval:]), ('eval', src[:n_eval], tgt[:n_eval])):
        for lang, sents in (('de', a), ('en', b)):
            with open(os.path.join(seg_dir, '%s.%s.bpe' % (name, lang)), 'w', encoding='utf-8') as f:
                f.write('\n'.join(pieces(s) for s in sents) + '\n')
    print('wrote %d train / %d eval pairs under %s' % (n - n_eval, n_eval, seg_dir))


if __name__ == '__main__':
    main()
