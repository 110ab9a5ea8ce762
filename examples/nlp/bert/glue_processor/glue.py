"""GLUE task readers: TSV files of a task directory -> (text_a, text_b, label) examples -> BERT
input arrays.

reference: examples/nlp/bert/glue_processor/glue.py (InputExample / InputFeatures, one processor
per task with get_train_examples / get_dev_examples / get_labels, convert_examples_to_features).
Tasks: CoLA, SST-2, MRPC, QQP, MNLI (matched), QNLI, RTE, WNLI (classification)."""
import csv
import os

import numpy as np


class InputExample(object):
    def __init__(self, guid, text_a, text_b=None, label=None):
        self.guid, self.text_a, self.text_b, self.label = guid, text_a, text_b, label


class DataProcessor(object):
    #: (column of text_a, column of text_b or None, column of the label, has header row) per split
    train_cols = dev_cols = None
    labels = ['0', '1']
    train_file, dev_file = 'train.tsv', 'dev.tsv'

    def get_labels(self):
        return list(self.labels)

    @staticmethod
    def _read_tsv(path):
        with open(path, encoding='utf-8') as f:
            return list(csv.reader(f, delimiter='\t', quotechar=None))

    def _examples(self, rows, cols, split):
        a, b, lab, header = cols
        out = []
        for i, row in enumerate(rows[1:] if header else rows):
            if max(a, b or 0, lab) >= len(row):
                continue
            out.append(InputExample('%s-%d' % (split, i), row[a], row[b] if b is not None else None, row[lab]))
        return out

    def get_train_examples(self, data_dir):
        return self._examples(self._read_tsv(os.path.join(data_dir, self.train_file)), self.train_cols, 'train')

    def get_dev_examples(self, data_dir):
        return self._examples(self._read_tsv(os.path.join(data_dir, self.dev_file)), self.dev_cols or self.train_cols, 'dev')


class ColaProcessor(DataProcessor):
    train_cols = (3, None, 1, False)


class Sst2Processor(DataProcessor):
    train_cols = (0, None, 1, True)


class MrpcProcessor(DataProcessor):
    train_cols = (3, 4, 0, True)


class QqpProcessor(DataProcessor):
    train_cols = (3, 4, 5, True)


class MnliProcessor(DataProcessor):
    train_cols = (8, 9, 11, True)
    dev_cols = (8, 9, 15, True)
    dev_file = 'dev_matched.tsv'
    labels = ['contradiction', 'entailment', 'neutral']


class QnliProcessor(DataProcessor):
    train_cols = (1, 2, 3, True)
    labels = ['entailment', 'not_entailment']


class RteProcessor(QnliProcessor):
    pass


class WnliProcessor(DataProcessor):
    train_cols = (1, 2, 3, True)


PROCESSORS = {'cola': ColaProcessor, 'sst-2': Sst2Processor, 'mrpc': MrpcProcessor, 'qqp': QqpProcessor,
              'mnli': MnliProcessor, 'qnli': QnliProcessor, 'rte': RteProcessor, 'wnli': WnliProcessor}


def convert_examples_to_features(examples, label_list, max_seq_length, tokenizer):
    """[CLS] a [SEP] (b [SEP]) padded to ``max_seq_length``; returns dict of int64 arrays."""
    label_map = {l: i for i, l in enumerate(label_list)}
    n = len(examples)
    ids = np.zeros((n, max_seq_length), dtype=np.int64)
    seg = np.zeros((n, max_seq_length), dtype=np.int64)
    att = np.zeros((n, max_seq_length), dtype=np.int64)
    lab = np.zeros(n, dtype=np.int64)
    for k, ex in enumerate(examples):
        a = tokenizer.tokenize(ex.text_a)
        b = tokenizer.tokenize(ex.text_b) if ex.text_b else None
        if b is not None:
            while len(a) + len(b) > max_seq_length - 3:
                (a if len(a) > len(b) else b).pop()
        else:
            a = a[:max_seq_length - 2]
        tokens = ['[CLS]'] + a + ['[SEP]']
        segment = [0] * len(tokens)
        if b is not None:
            tokens += b + ['[SEP]']
            segment += [1] * (len(b) + 1)
        t = tokenizer.convert_tokens_to_ids(tokens)
        ids[k, :len(t)], seg[k, :len(t)], att[k, :len(t)] = t, segment, 1
        lab[k] = label_map[ex.label]
    return dict(input_ids=ids, token_type_ids=seg, attention_mask=att, labels=lab)
