"""BERT WordPiece tokenizer.  reference: python/hetu/tokenizers/bert_tokenizer.py:76-363
(same behaviour as the original BERT ``tokenization.py``: basic whitespace /
punctuation / CJK splitting, optional lower-casing + accent stripping, greedy
longest-match-first WordPiece)."""
from __future__ import annotations

import collections
import unicodedata
from typing import Dict, Iterable, List


def load_vocab(vocab_file: str) -> "collections.OrderedDict[str, int]":
    vocab = collections.OrderedDict()
    with open(vocab_file, 'r', encoding='utf-8') as reader:
        for index, token in enumerate(reader):
            vocab[token.rstrip('\n')] = index
    return vocab


def whitespace_tokenize(text: str) -> List[str]:
    text = text.strip()
    return text.split() if text else []


def _is_whitespace(ch):
    return ch in ' \t\n\r' or unicodedata.category(ch) == 'Zs'


def _is_control(ch):
    if ch in '\t\n\r':
        return False
    return unicodedata.category(ch) in ('Cc', 'Cf')


def _is_punctuation(ch):
    cp = ord(ch)
    if 33 <= cp <= 47 or 58 <= cp <= 64 or 91 <= cp <= 96 or 123 <= cp <= 126:
        return True
    return unicodedata.category(ch).startswith('P')


def _is_cjk(cp):
    return (0x4E00 <= cp <= 0x9FFF or 0x3400 <= cp <= 0x4DBF or 0x20000 <= cp <= 0x2A6DF
            or 0x2A700 <= cp <= 0x2B73F or 0x2B740 <= cp <= 0x2B81F or 0x2B820 <= cp <= 0x2CEAF
            or 0xF900 <= cp <= 0xFAFF or 0x2F800 <= cp <= 0x2FA1F)


class BasicTokenizer(object):
    def __init__(self, do_lower_case=True, never_split=("[UNK]", "[SEP]", "[PAD]", "[CLS]", "[MASK]")):
        self.do_lower_case = do_lower_case
        self.never_split = set(never_split)

    def tokenize(self, text: str) -> List[str]:
        out = []
        for ch in text:
            cp = ord(ch)
            if cp == 0 or cp == 0xFFFD or _is_control(ch):
                continue
            if _is_whitespace(ch):
                out.append(' ')
            elif _is_cjk(cp):
                out.extend([' ', ch, ' '])
            else:
                out.append(ch)
        tokens = []
        for tok in whitespace_tokenize(''.join(out)):
            if tok in self.never_split:
                tokens.append(tok)
                continue
            if self.do_lower_case:
                tok = ''.join(c for c in unicodedata.normalize('NFD', tok.lower())
                              if unicodedata.category(c) != 'Mn')
            cur = []
            for ch in tok:
                if _is_punctuation(ch):
                    if cur:
                        tokens.append(''.join(cur))
                        cur = []
                    tokens.append(ch)
                else:
                    cur.append(ch)
            if cur:
                tokens.append(''.join(cur))
        return tokens


class WordpieceTokenizer(object):
    def __init__(self, vocab: Dict[str, int], unk_token="[UNK]", max_input_chars_per_word=100):
        self.vocab, self.unk_token = vocab, unk_token
        self.max_input_chars_per_word = max_input_chars_per_word

    def tokenize(self, text: str) -> List[str]:
        out = []
        for token in whitespace_tokenize(text):
            if len(token) > self.max_input_chars_per_word:
                out.append(self.unk_token)
                continue
            start, pieces, bad = 0, [], False
            while start < len(token):
                end, cur = len(token), None
                while start < end:
                    sub = token[start:end]
                    if start > 0:
                        sub = '##' + sub
                    if sub in self.vocab:
                        cur = sub
                        break
                    end -= 1
                if cur is None:
                    bad = True
                    break
                pieces.append(cur)
                start = end
            out.extend([self.unk_token] if bad else pieces)
        return out


class BertTokenizer(object):
    def __init__(self, vocab_file=None, vocab=None, do_lower_case=True, max_len=None,
                 never_split=("[UNK]", "[SEP]", "[PAD]", "[CLS]", "[MASK]")):
        if vocab is None:
            vocab = load_vocab(vocab_file)
        elif not isinstance(vocab, dict):
            vocab = collections.OrderedDict((t, i) for i, t in enumerate(vocab))
        self.vocab = vocab
        self.ids_to_tokens = collections.OrderedDict((i, t) for t, i in vocab.items())
        self.basic_tokenizer = BasicTokenizer(do_lower_case, never_split)
        self.wordpiece_tokenizer = WordpieceTokenizer(self.vocab)
        self.max_len = max_len if max_len is not None else int(1e12)

    def tokenize(self, text: str) -> List[str]:
        out = []
        for tok in self.basic_tokenizer.tokenize(text):
            out.extend(self.wordpiece_tokenizer.tokenize(tok))
        return out

    def convert_tokens_to_ids(self, tokens: Iterable[str]) -> List[int]:
        unk = self.vocab.get('[UNK]', 0)
        ids = [self.vocab.get(t, unk) for t in tokens]
        if len(ids) > self.max_len:
            raise ValueError('Token indices sequence length is longer than the maximum (%d > %d)'
                             % (len(ids), self.max_len))
        return ids

    def convert_ids_to_tokens(self, ids: Iterable[int]) -> List[str]:
        return [self.ids_to_tokens[int(i)] for i in ids]

    def encode(self, text_a, text_b=None, max_seq_length=128):
        """-> (input_ids, token_type_ids, attention_mask) padded to ``max_seq_length``."""
        ta = self.tokenize(text_a)
        tb = self.tokenize(text_b) if text_b is not None else []
        budget = max_seq_length - (3 if tb else 2)
        while len(ta) + len(tb) > budget:
            (ta if len(ta) >= len(tb) else tb).pop()
        tokens = ['[CLS]'] + ta + ['[SEP]']
        types = [0] * len(tokens)
        if tb:
            tokens += tb + ['[SEP]']
            types += [1] * (len(tb) + 1)
        ids = self.convert_tokens_to_ids(tokens)
        mask = [1] * len(ids)
        pad = max_seq_length - len(ids)
        pid = self.vocab.get('[PAD]', 0)
        return ids + [pid] * pad, types + [0] * pad, mask + [0] * pad
