#!/usr/bin/env python
"""Turn raw text into BERT pre-training instances (masked LM + next-sentence prediction).

reference: examples/nlp/bert/create_pretraining_data.py (the original BERT recipe: documents of
sentences -> sentence pairs A/B with 50 % random B, 15 % of the tokens masked: 80 % [MASK], 10 %
random, 10 % kept; at most ``max_predictions_per_seq``) and ``load_data.py`` (the array layout the
trainers read).  Input: one sentence per line, blank line between documents.  Output: one ``.npz``
with input_ids, token_type_ids, attention_mask, masked_lm_labels (-1 = not predicted),
next_sentence_label - the file ``train_hetu_bert.py --data`` takes.

    python create_pretraining_data.py --input corpus.txt --output train.npz --vocab vocab.txt
    python create_pretraining_data.py --input corpus.txt --output train.npz --build-vocab vocab.txt
"""
import argparse
import collections
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', '..', '..'))
from hetu_61a7_b200.utils.tokenizers import BasicTokenizer, BertTokenizer   # noqa: E402

SPECIAL = ['[PAD]', '[UNK]', '[CLS]', '[SEP]', '[MASK]']


def read_documents(path):
    docs, cur = [], []
    with open(path, encoding='utf-8') as f:
        for line in f:
            line = line.strip()
            if not line:
                if cur:
                    docs.append(cur)
                    cur = []
            else:
                cur.append(line)
    if cur:
        docs.append(cur)
    return docs


def build_vocab(docs, path, max_size=30522, min_count=1):
    """Whole-word vocabulary from the corpus (enough for experiments; real runs use the published
    wordpiece vocab)."""
    tok = BasicTokenizer(do_lower_case=True)
    cnt = collections.Counter(w for d in docs for s in d for w in tok.tokenize(s))
    words = [w for w, c in cnt.most_common(max_size - len(SPECIAL)) if c >= min_count]
    with open(path, 'w', encoding='utf-8') as f:
        f.write('\n'.join(SPECIAL + words) + '\n')
    return path


def truncate_pair(a, b, max_tokens, rng):
    while len(a) + len(b) > max_tokens:
        longer = a if len(a) > len(b) else b
        if rng.rand() < 0.5:
            del longer[0]
        else:
            longer.pop()


def mask_tokens(tokens, vocab_words, masked_lm_prob, max_predictions, rng):
    cand = [i for i, t in enumerate(tokens) if t not in ('[CLS]', '[SEP]')]
    rng.shuffle(cand)
    n_pred = min(max_predictions, max(1, int(round(len(tokens) * masked_lm_prob))))
    out = list(tokens)
    labels = {}
    for i in sorted(cand[:n_pred]):
        r = rng.rand()
        labels[i] = tokens[i]
        out[i] = '[MASK]' if r < 0.8 else (vocab_words[rng.randint(len(vocab_words))] if r < 0.9 else tokens[i])
    return out, labels


def create_instances(docs_tok, max_seq_length, dupe_factor, short_seq_prob, masked_lm_prob, max_predictions, vocab_words, rng):
    instances = []
    max_tokens = max_seq_length - 3
    for _ in range(dupe_factor):
        for di, doc in enumerate(docs_tok):
            target = max_tokens if rng.rand() >= short_seq_prob else rng.randint(2, max_tokens + 1)
            chunk, length, i = [], 0, 0
            while i < len(doc):
                chunk.append(doc[i])
                length += len(doc[i])
                if i == len(doc) - 1 or length >= target:
                    if chunk:
                        a_end = 1 if len(chunk) < 2 else rng.randint(1, len(chunk))
                        a = [t for s in chunk[:a_end] for t in s]
                        b, is_random = [], False
                        if len(chunk) == 1 or (len(docs_tok) > 1 and rng.rand() < 0.5):
                            is_random = True
                            for _try in range(10):
                                rd = rng.randint(len(docs_tok))
                                if rd != di:
                                    break
                            rdoc = docs_tok[rd]
                            start = rng.randint(len(rdoc))
                            for s in rdoc[start:]:
                                b.extend(s)
                                if len(b) >= target - len(a):
                                    break
                            i -= len(chunk) - a_end              # put the unused sentences back
                        else:
                            b = [t for s in chunk[a_end:] for t in s]
                        a, b = list(a), list(b)
                        if a and b:
                            truncate_pair(a, b, max_tokens, rng)
                            tokens = ['[CLS]'] + a + ['[SEP]'] + b + ['[SEP]']
                            seg = [0] * (len(a) + 2) + [1] * (len(b) + 1)
                            masked, labels = mask_tokens(tokens, vocab_words, masked_lm_prob, max_predictions, rng)
                            instances.append((masked, seg, labels, int(is_random)))
                    chunk, length = [], 0
                    target = max_tokens if rng.rand() >= short_seq_prob else rng.randint(2, max_tokens + 1)
                i += 1
    rng.shuffle(instances)
    return instances


def to_arrays(instances, tokenizer, max_seq_length):
    n = len(instances)
    ids = np.zeros((n, max_seq_length), dtype=np.int64)
    seg = np.zeros((n, max_seq_length), dtype=np.int64)
    att = np.zeros((n, max_seq_length), dtype=np.int64)
    lab = np.full((n, max_seq_length), -1, dtype=np.int64)
    nsp = np.zeros(n, dtype=np.int64)
    for k, (tokens, segment, labels, is_random) in enumerate(instances):
        t = tokenizer.convert_tokens_to_ids(tokens)
        ids[k, :len(t)] = t
        seg[k, :len(t)] = segment
        att[k, :len(t)] = 1
        for pos, word in labels.items():
            lab[k, pos] = tokenizer.convert_tokens_to_ids([word])[0]
        nsp[k] = is_random
    return dict(input_ids=ids, token_type_ids=seg, attention_mask=att, masked_lm_labels=lab, next_sentence_label=nsp)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--input', required=True)
    ap.add_argument('--output', required=True)
    ap.add_argument('--vocab', default=None)
    ap.add_argument('--build-vocab', default=None, help='write a whole-word vocabulary built from the corpus here and use it')
    ap.add_argument('--max-seq-length', type=int, default=128)
    ap.add_argument('--max-predictions-per-seq', type=int, default=20)
    ap.add_argument('--masked-lm-prob', type=float, default=0.15)
    ap.add_argument('--short-seq-prob', type=float, default=0.1)
    ap.add_argument('--dupe-factor', type=int, default=5)
    ap.add_argument('--random-seed', type=int, default=12345)
    args = ap.parse_args()
    docs = read_documents(args.input)
    vocab_file = args.vocab or build_vocab(docs, args.build_vocab or args.output + '.vocab.txt')
    tokenizer = BertTokenizer(vocab_file=vocab_file, do_lower_case=True)
    docs_tok = [[tokenizer.tokenize(s) for s in d] for d in docs]
    docs_tok = [[s for s in d if s] for d in docs_tok]
    docs_tok = [d for d in docs_tok if d]
    rng = np.random.RandomState(args.random_seed)
    vocab_words = [w for w in tokenizer.vocab if w not in SPECIAL]
    inst = create_instances(docs_tok, args.max_seq_length, args.dupe_factor, args.short_seq_prob, args.masked_lm_prob,
                            args.max_predictions_per_seq, vocab_words, rng)
    arrays = to_arrays(inst, tokenizer, args.max_seq_length)
    np.savez(args.output, **arrays)
    print('wrote %d instances (%d documents, vocab %d) to %s' % (len(inst), len(docs_tok), len(tokenizer.vocab), args.output))


if __name__ == '__main__':
    main()
