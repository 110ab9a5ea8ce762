"""NLP models: BERT (base / large, MLM + NSP pre-training heads, sequence classification) and an
encoder-decoder Transformer.

reference: examples/nlp/bert/hetu_bert.py:21-942 (BertEmbeddings / BertLayer / BertModel /
BertForPreTraining / BertForSequenceClassification, config examples/nlp/bert/bert_config.py) and
examples/nlp/hetu_transformer.py:1-266.  Hidden states are kept flat as [batch*seq, hidden]
so every projection is one 2-D GEMM on the tcgen05 kernel.  ``fused_attention=True`` (default)
uses ``ht.fused_attention_op``; ``False`` builds the reference's five-op attention
(batch_matmul, mask add, softmax, dropout, batch_matmul) for parity checks.
"""
from __future__ import annotations

import math

import numpy as np

import hetu as ht
from hetu import init


class BertConfig(object):
    """reference bert_config.py:1-60."""

    def __init__(self, vocab_size=30522, hidden_size=768, num_hidden_layers=12, num_attention_heads=12,
                 intermediate_size=3072, hidden_act='gelu', hidden_dropout_prob=0.1,
                 attention_probs_dropout_prob=0.1, max_position_embeddings=512, type_vocab_size=2,
                 initializer_range=0.02, batch_size=32, seq_len=128, fused_attention=True, tp=1,
                 tie_word_embeddings=True):
        self.vocab_size, self.hidden_size = vocab_size, hidden_size
        self.num_hidden_layers, self.num_attention_heads = num_hidden_layers, num_attention_heads
        self.intermediate_size, self.hidden_act = intermediate_size, hidden_act
        self.hidden_dropout_prob = hidden_dropout_prob
        self.attention_probs_dropout_prob = attention_probs_dropout_prob
        self.max_position_embeddings, self.type_vocab_size = max_position_embeddings, type_vocab_size
        self.initializer_range = initializer_range
        self.batch_size, self.seq_len = batch_size, seq_len
        self.fused_attention = fused_attention
        #: tensor-parallel degree of the feed-forward block (build the model inside a tuple context;
        #: w1 is column-parallel, w2 row-parallel, Megatron style) - see parallel/mp_lowering.py
        self.tp = tp
        #: the MLM decoder shares the word-embedding matrix (reference hetu_bert.py); pipeline-parallel
        #: layouts untie it so the table does not travel between the first and the last stage
        self.tie_word_embeddings = tie_word_embeddings

    @classmethod
    def base(cls, **kw):
        return cls(**kw)

    @classmethod
    def large(cls, **kw):
        """BERT-large (scripts/train_hetu_bert_large_dp.sh: hidden 1024, 24 layers, 16 heads)."""
        d = dict(hidden_size=1024, num_hidden_layers=24, num_attention_heads=16, intermediate_size=4096)
        d.update(kw)
        return cls(**d)


def _dense(x, in_dim, out_dim, name, std, act=None):
    w = init.truncated_normal((in_dim, out_dim), stddev=std, name=name + '_weight')
    b = init.zeros((out_dim,), name=name + '_bias')
    y = ht.linear_op(x, w, b)
    if act == 'gelu':
        y = ht.gelu_op(y)
    elif act == 'relu':
        y = ht.relu_op(y)
    elif act == 'tanh':
        y = ht.tanh_op(y)
    return y


def _layer_norm(x, dim, name, eps=1e-12):
    return ht.layer_normalization_op(x, init.ones((dim,), name=name + '_scale'),
                                     init.zeros((dim,), name=name + '_bias'), eps=eps)


def _dropout(x, p):
    return ht.dropout_op(x, 1.0 - p) if p > 0 else x


def _split_heads(x, batch, seq, heads, head_dim):
    x = ht.array_reshape_op(x, [batch, seq, heads, head_dim])
    return ht.transpose_op(x, [0, 2, 1, 3])                      # [B, H, S, D]


def attention(q_in, kv_in, mask, cfg_or_dims, name, std, batch, seq_q, seq_k, fused=True, attn_dropout=0.0):
    """Multi-head attention on flat [B*S, hidden] states.  ``mask``: additive [B, 1, 1, S_k] or None."""
    hidden, heads = cfg_or_dims
    hd = hidden // heads
    q = _split_heads(_dense(q_in, hidden, hidden, name + '_query', std), batch, seq_q, heads, hd)
    k = _split_heads(_dense(kv_in, hidden, hidden, name + '_key', std), batch, seq_k, heads, hd)
    v = _split_heads(_dense(kv_in, hidden, hidden, name + '_value', std), batch, seq_k, heads, hd)
    if fused and attn_dropout == 0.0:
        ctx_layer = ht.fused_attention_op(q, k, v, mask)
    else:
        scores = ht.batch_matmul_op(q, k, trans_B=True)
        scores = ht.mul_byconst_op(scores, 1.0 / math.sqrt(hd))
        if mask is not None:
            scores = scores + ht.broadcastto_op(mask, scores)
        probs = _dropout(ht.softmax_op(scores), attn_dropout)
        ctx_layer = ht.batch_matmul_op(probs, v)
    ctx_layer = ht.transpose_op(ctx_layer, [0, 2, 1, 3])
    ctx_layer = ht.array_reshape_op(ctx_layer, [batch * seq_q, hidden])
    return _dense(ctx_layer, hidden, hidden, name + '_output', std)


class BertModel(object):
    """Embeddings + encoder stack + pooler.  ``__call__(input_ids, token_type_ids, attention_mask)``
    returns (sequence_output [B*S, H], pooled_output [B, H])."""

    def __init__(self, config: BertConfig, name='bert', stage_contexts=None):
        """``stage_contexts``: list of contexts (anything ``ht.context`` accepts), one per pipeline stage;
        the encoder layers are split evenly, embeddings go to the first and the pooler to the last."""
        self.cfg, self.name = config, name
        self.stage_contexts = list(stage_contexts) if stage_contexts else None

    def _stage(self, i):
        import contextlib
        if not self.stage_contexts:
            return contextlib.nullcontext()
        S, L = len(self.stage_contexts), self.cfg.num_hidden_layers
        return ht.context(self.stage_contexts[min(S - 1, i * S // max(L, 1))] if i >= 0 else self.stage_contexts[0])

    def embeddings(self, input_ids, token_type_ids):
        c = self.cfg
        std = c.initializer_range
        word = init.truncated_normal((c.vocab_size, c.hidden_size), stddev=std, name=self.name + '_word_embeddings')
        pos = init.truncated_normal((c.max_position_embeddings, c.hidden_size), stddev=std,
                                    name=self.name + '_position_embeddings')
        typ = init.truncated_normal((c.type_vocab_size, c.hidden_size), stddev=std,
                                    name=self.name + '_token_type_embeddings')
        self.word_embeddings = word
        pos_ids = ht.Variable(self.name + '_position_ids', trainable=False,
                              value=np.tile(np.arange(c.seq_len, dtype=np.float32), (c.batch_size, 1)))
        x = ht.embedding_lookup_op(word, input_ids) + ht.embedding_lookup_op(pos, pos_ids) \
            + ht.embedding_lookup_op(typ, token_type_ids)
        x = ht.array_reshape_op(x, [c.batch_size * c.seq_len, c.hidden_size])
        return _dropout(_layer_norm(x, c.hidden_size, self.name + '_emb_ln'), c.hidden_dropout_prob)

    def layer(self, x, mask, i):
        c = self.cfg
        std, n = c.initializer_range, '%s_layer%d' % (self.name, i)
        a = attention(x, x, mask, (c.hidden_size, c.num_attention_heads), n + '_attn', std, c.batch_size,
                      c.seq_len, c.seq_len, fused=c.fused_attention, attn_dropout=c.attention_probs_dropout_prob
                      if not c.fused_attention else 0.0)
        x = _layer_norm(x + _dropout(a, c.hidden_dropout_prob), c.hidden_size, n + '_attn_ln')
        if c.tp > 1:
            h = self._ffn_tp(x, n, std)
        else:
            h = _dense(x, c.hidden_size, c.intermediate_size, n + '_intermediate', std, act=c.hidden_act)
            h = _dense(h, c.intermediate_size, c.hidden_size, n + '_ffn_output', std)
        return _layer_norm(x + _dropout(h, c.hidden_dropout_prob), c.hidden_size, n + '_ffn_ln')

    def _ffn_tp(self, x, n, std):
        """Column-parallel w1 / row-parallel w2: one all-reduce of the block output per direction."""
        c = self.cfg
        t = c.tp
        w1 = init.truncated_normal((c.hidden_size, c.intermediate_size), stddev=std, name=n + '_intermediate_weight')
        b1 = init.zeros((c.intermediate_size,), name=n + '_intermediate_bias')
        w2 = init.truncated_normal((c.intermediate_size, c.hidden_size), stddev=std, name=n + '_ffn_output_weight')
        b2 = init.zeros((c.hidden_size,), name=n + '_ffn_output_bias')
        h = ht.matmul_op(x, ht.dispatch(w1, (1, t)))
        h = h + ht.broadcastto_op(ht.dispatch(b1, (t,)), h)
        h = ht.gelu_op(h) if c.hidden_act == 'gelu' else ht.relu_op(h)
        out = ht.dispatch(ht.matmul_op(h, ht.dispatch(w2, (t, 1))), (1, 1))        # partial sums -> all-reduce
        return out + ht.broadcastto_op(b2, out)

    def __call__(self, input_ids, token_type_ids, attention_mask=None):
        c = self.cfg
        mask = None
        if attention_mask is not None:                       # 1 = attend, 0 = pad  ->  additive -10000
            m = ht.array_reshape_op(attention_mask, [c.batch_size, 1, 1, c.seq_len])
            mask = ht.mul_byconst_op(ht.addbyconst_op(m, -1.0), 10000.0)
        with self._stage(-1):
            x = self.embeddings(input_ids, token_type_ids)
        for i in range(c.num_hidden_layers):
            with self._stage(i):
                x = self.layer(x, mask, i)
        if self.stage_contexts:
            with ht.context(self.stage_contexts[-1]):
                return x, self._pool(x)
        return x, self._pool(x)

    def _pool(self, x):
        c = self.cfg
        first = ht.slice_op(ht.array_reshape_op(x, [c.batch_size, c.seq_len, c.hidden_size]), (0, 0, 0),
                            (-1, 1, -1))
        first = ht.array_reshape_op(first, [c.batch_size, c.hidden_size])
        return _dense(first, c.hidden_size, c.hidden_size, self.name + '_pooler', c.initializer_range, act='tanh')


class BertForPreTraining(object):
    """MLM (decoder tied to the word embeddings) + NSP heads.  Returns (mlm_logits, nsp_logits,
    mlm_loss, nsp_loss); total loss = mlm_loss + nsp_loss (reference hetu_bert.py:618-660)."""

    def __init__(self, config: BertConfig, stage_contexts=None):
        self.cfg = config
        self.bert = BertModel(config, stage_contexts=stage_contexts)

    def __call__(self, input_ids, token_type_ids, attention_mask, masked_lm_labels=None, next_sentence_label=None):
        import contextlib
        sc = self.bert.stage_contexts
        with (ht.context(sc[-1]) if sc else contextlib.nullcontext()):
            return self._build(input_ids, token_type_ids, attention_mask, masked_lm_labels, next_sentence_label)

    def _build(self, input_ids, token_type_ids, attention_mask, masked_lm_labels, next_sentence_label):
        c = self.cfg
        seq, pooled = self.bert(input_ids, token_type_ids, attention_mask)
        h = _dense(seq, c.hidden_size, c.hidden_size, 'cls_transform', c.initializer_range, act=c.hidden_act)
        h = _layer_norm(h, c.hidden_size, 'cls_transform_ln')
        bias = init.zeros((c.vocab_size,), name='cls_decoder_bias')
        decoder = self.bert.word_embeddings if c.tie_word_embeddings else \
            init.truncated_normal((c.vocab_size, c.hidden_size), stddev=c.initializer_range, name='cls_decoder_weight')
        mlm_logits = ht.linear_op(h, decoder, bias, trans_B=True)                        # [B*S, vocab]
        nsp_logits = _dense(pooled, c.hidden_size, 2, 'cls_seq_relationship', c.initializer_range)
        out = [mlm_logits, nsp_logits]
        if masked_lm_labels is not None and next_sentence_label is not None:
            labels = ht.array_reshape_op(masked_lm_labels, [c.batch_size * c.seq_len])
            mlm = ht.softmaxcrossentropy_sparse_op(mlm_logits, labels, ignored_index=-1)
            nsp = ht.softmaxcrossentropy_sparse_op(nsp_logits, next_sentence_label, ignored_index=-1)
            out += [ht.reduce_mean_op(mlm, [0]), ht.reduce_mean_op(nsp, [0])]
        return out


class BertForSequenceClassification(object):
    """GLUE fine-tuning head (reference hetu_bert.py:700-760)."""

    def __init__(self, config: BertConfig, num_labels=2):
        self.cfg, self.num_labels = config, num_labels
        self.bert = BertModel(config)

    def __call__(self, input_ids, token_type_ids, attention_mask, labels=None):
        c = self.cfg
        _, pooled = self.bert(input_ids, token_type_ids, attention_mask)
        logits = _dense(_dropout(pooled, c.hidden_dropout_prob), c.hidden_size, self.num_labels, 'classifier',
                        c.initializer_range)
        if labels is None:
            return logits
        loss = ht.reduce_mean_op(ht.softmaxcrossentropy_sparse_op(logits, labels), [0])
        return logits, loss


# ------------------------------------------------------------------------------ transformer
class TransformerHParams(object):
    """reference examples/nlp/hparams.py (defaults of the IWSLT de-en recipe)."""

    def __init__(self, vocab_size=32000, d_model=512, d_ff=2048, num_blocks=6, num_heads=8, maxlen1=100,
                 maxlen2=100, dropout_rate=0.3, batch_size=16):
        self.vocab_size, self.d_model, self.d_ff = vocab_size, d_model, d_ff
        self.num_blocks, self.num_heads = num_blocks, num_heads
        self.maxlen1, self.maxlen2, self.dropout_rate, self.batch_size = maxlen1, maxlen2, dropout_rate, batch_size


def _positional_encoding(batch, maxlen, d_model, name):
    pe = np.array([[pos / np.power(10000, (i - i % 2) / d_model) for i in range(d_model)]
                   for pos in range(maxlen)], dtype=np.float32)
    pe[:, 0::2] = np.sin(pe[:, 0::2])
    pe[:, 1::2] = np.cos(pe[:, 1::2])
    return ht.Variable(name, value=np.tile(pe[None], (batch, 1, 1)).reshape(batch * maxlen, d_model),
                       trainable=False)


class Transformer(object):
    """Encoder-decoder Transformer with shared source/target embedding; ``__call__(xs, ys_in, ys_out)``
    with float-encoded id matrices [B, T]; id 0 is padding.  Returns (loss, logits)."""

    def __init__(self, hp: TransformerHParams):
        self.hp = hp
        self.embeddings = init.xavier_normal((hp.vocab_size, hp.d_model), name='transformer_embed')

    def _embed(self, ids, maxlen, name):
        hp = self.hp
        x = ht.embedding_lookup_op(self.embeddings, ids)
        x = ht.array_reshape_op(x, [hp.batch_size * maxlen, hp.d_model])
        x = ht.mul_byconst_op(x, hp.d_model ** 0.5) + _positional_encoding(hp.batch_size, maxlen, hp.d_model, name)
        return _dropout(x, hp.dropout_rate)

    def _ffn(self, x, name):
        hp = self.hp
        h = _dense(x, hp.d_model, hp.d_ff, name + '_ff1', 0.02, act='relu')
        h = _dense(h, hp.d_ff, hp.d_model, name + '_ff2', 0.02)
        return _layer_norm(x + h, hp.d_model, name + '_ffn_ln', eps=1e-8)

    def encode(self, xs, src_mask):
        hp = self.hp
        x = self._embed(xs, hp.maxlen1, 'transformer_pe_src')
        for i in range(hp.num_blocks):
            n = 'enc%d' % i
            a = attention(x, x, src_mask, (hp.d_model, hp.num_heads), n + '_self', 0.02, hp.batch_size,
                          hp.maxlen1, hp.maxlen1)
            x = _layer_norm(x + _dropout(a, hp.dropout_rate), hp.d_model, n + '_self_ln', eps=1e-8)
            x = self._ffn(x, n)
        return x

    def decode(self, ys, memory, src_mask, causal_mask):
        hp = self.hp
        x = self._embed(ys, hp.maxlen2, 'transformer_pe_tgt')
        for i in range(hp.num_blocks):
            n = 'dec%d' % i
            a = attention(x, x, causal_mask, (hp.d_model, hp.num_heads), n + '_self', 0.02, hp.batch_size,
                          hp.maxlen2, hp.maxlen2, fused=False)           # [S, S] mask: unfused path
            x = _layer_norm(x + _dropout(a, hp.dropout_rate), hp.d_model, n + '_self_ln', eps=1e-8)
            a = attention(x, memory, src_mask, (hp.d_model, hp.num_heads), n + '_cross', 0.02, hp.batch_size,
                          hp.maxlen2, hp.maxlen1)
            x = _layer_norm(x + _dropout(a, hp.dropout_rate), hp.d_model, n + '_cross_ln', eps=1e-8)
            x = self._ffn(x, n)
        return x

    def __call__(self, xs, ys_in, ys_out, src_mask=None):
        """``src_mask``: additive [B, 1, 1, maxlen1] key-padding mask (0 / -1e9), optional."""
        hp = self.hp
        causal = np.triu(np.full((hp.maxlen2, hp.maxlen2), -1e9, dtype=np.float32), k=1)
        causal_mask = ht.Variable('transformer_causal_mask', trainable=False,
                                  value=np.tile(causal[None, None], (hp.batch_size, 1, 1, 1)))
        memory = self.encode(xs, src_mask)
        dec = self.decode(ys_in, memory, src_mask, causal_mask)
        logits = ht.matmul_op(dec, self.embeddings, trans_B=True)        # tied output projection
        labels = ht.array_reshape_op(ys_out, [hp.batch_size * hp.maxlen2])
        loss = ht.softmaxcrossentropy_sparse_op(logits, labels, ignored_index=0)
        return ht.reduce_mean_op(loss, [0]), logits
