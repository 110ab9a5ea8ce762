"""Parallel-corpus loading for the Transformer example (reference examples/nlp/data_load.py:
load_vocab, load_data (length filter), encode (<s>, </s>, <unk>), fixed-length padded batches).
Vocabulary convention: id 0 <pad>, 1 <unk>, 2 <s>, 3 </s>."""
import numpy as np


def load_vocab(vocab_fpath):
    vocab = [line.split()[0] for line in open(vocab_fpath, encoding='utf-8').read().splitlines() if line.strip()]
    token2idx = {tok: idx for idx, tok in enumerate(vocab)}
    idx2token = {idx: tok for idx, tok in enumerate(vocab)}
    return token2idx, idx2token


def load_data(fpath1, fpath2, maxlen1, maxlen2):
    sents1, sents2 = [], []
    with open(fpath1, encoding='utf-8') as f1, open(fpath2, encoding='utf-8') as f2:
        for s1, s2 in zip(f1, f2):
            if len(s1.split()) + 1 > maxlen1 or len(s2.split()) + 1 > maxlen2:
                continue
            sents1.append(s1.strip())
            sents2.append(s2.strip())
    return sents1, sents2


def encode(sent, kind, token2idx):
    tokens = sent.split() + ['</s>'] if kind == 'x' else ['<s>'] + sent.split() + ['</s>']
    unk = token2idx.get('<unk>', 1)
    return [token2idx.get(t, unk) for t in tokens]


class DataLoader(object):
    """Padded id matrices: xs [N, maxlen1], decoder inputs [N, maxlen2] (starts with <s>) and
    decoder targets [N, maxlen2] (ends with </s>)."""

    def __init__(self, fpath1, fpath2, maxlen1, maxlen2, vocab_fpath):
        self.token2idx, self.idx2token = load_vocab(vocab_fpath)
        s1, s2 = load_data(fpath1, fpath2, maxlen1, maxlen2)
        n = len(s1)
        self.xs = np.zeros((n, maxlen1), dtype=np.float32)
        self.ys_in = np.zeros((n, maxlen2), dtype=np.float32)
        self.ys_out = np.zeros((n, maxlen2), dtype=np.float32)
        for k, (a, b) in enumerate(zip(s1, s2)):
            x, y = encode(a, 'x', self.token2idx), encode(b, 'y', self.token2idx)
            self.xs[k, :len(x)] = x
            self.ys_in[k, :len(y) - 1] = y[:-1]
            self.ys_out[k, :len(y) - 1] = y[1:]
        self.n = n

    def batches(self, batch_size, shuffle=True, seed=0):
        order = np.random.RandomState(seed).permutation(self.n) if shuffle else np.arange(self.n)
        for b in range(0, self.n - batch_size + 1, batch_size):
            idx = order[b:b + batch_size]
            yield self.xs[idx], self.ys_in[idx], self.ys_out[idx]
