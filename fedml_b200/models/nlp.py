"""Sequence models.

* ``RNN_OriginalFedAvg`` / ``RNN_StackOverFlow`` — reference ``model/nlp/rnn.py:5-86`` (key names
  ``embeddings/lstm/fc`` and ``word_embeddings/lstm/fc1/fc2``).
* ``DistilBertForSequenceClassification`` — the north-star FedNLP model, absent from the
  reference.  Written out explicitly (rather than wrapping ``transformers``) so the
  linear layers / attention / LayerNorm can be swapped for the grouped sm_100a kernels; parameter
  names follow the HuggingFace checkpoint layout so ``state_dict``s interchange
  (104 entries, 66,968,852 parameters with 20 labels).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class RNN_OriginalFedAvg(nn.Module):  # noqa: N801
    def __init__(self, embedding_dim=8, vocab_size=90, hidden_size=256):
        super().__init__()
        self.embeddings = nn.Embedding(num_embeddings=vocab_size, embedding_dim=embedding_dim, padding_idx=0)
        self.lstm = nn.LSTM(input_size=embedding_dim, hidden_size=hidden_size, num_layers=2, batch_first=True)
        self.fc = nn.Linear(hidden_size, vocab_size)

    def forward(self, input_seq):
        out, _ = self.lstm(self.embeddings(input_seq))
        return self.fc(out[:, -1])


class RNN_StackOverFlow(nn.Module):  # noqa: N801
    def __init__(self, vocab_size=10000, num_oov_buckets=1, embedding_size=96, latent_size=670, num_layers=1):
        super().__init__()
        ext = vocab_size + 3 + num_oov_buckets
        self.word_embeddings = nn.Embedding(num_embeddings=ext, embedding_dim=embedding_size, padding_idx=0)
        self.lstm = nn.LSTM(input_size=embedding_size, hidden_size=latent_size, num_layers=num_layers, batch_first=True)
        self.fc1 = nn.Linear(latent_size, embedding_size)
        self.fc2 = nn.Linear(embedding_size, ext)

    def forward(self, input_seq, hidden_state=None):
        out, hidden_state = self.lstm(self.word_embeddings(input_seq), hidden_state)
        logits = self.fc2(self.fc1(out))
        return logits.transpose(1, 2)        # [B, V, S] for CrossEntropyLoss(ignore_index=0)


# ------------------------------------------------------------------ DistilBERT
class _Embeddings(nn.Module):
    def __init__(self, vocab, dim, max_pos, dropout):
        super().__init__()
        self.word_embeddings = nn.Embedding(vocab, dim, padding_idx=0)
        self.position_embeddings = nn.Embedding(max_pos, dim)
        self.LayerNorm = nn.LayerNorm(dim, eps=1e-12)
        self.dropout = nn.Dropout(dropout)

    def forward(self, ids):
        pos = torch.arange(ids.shape[1], device=ids.device).unsqueeze(0)
        return self.dropout(self.LayerNorm(self.word_embeddings(ids) + self.position_embeddings(pos)))


class _SelfAttention(nn.Module):
    def __init__(self, dim, heads, dropout):
        super().__init__()
        self.n_heads, self.dim, self.p = heads, dim, dropout
        self.q_lin, self.k_lin = nn.Linear(dim, dim), nn.Linear(dim, dim)
        self.v_lin, self.out_lin = nn.Linear(dim, dim), nn.Linear(dim, dim)

    def forward(self, x, mask):
        B, S, D = x.shape
        h = self.n_heads
        split = lambda t: t.view(B, S, h, D // h).transpose(1, 2)
        q, k, v = split(self.q_lin(x)), split(self.k_lin(x)), split(self.v_lin(x))
        attn_mask = None if mask is None else mask[:, None, None, :].to(torch.bool)
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=attn_mask,
                                           dropout_p=self.p if self.training else 0.0)
        return self.out_lin(o.transpose(1, 2).reshape(B, S, D))


class _FFN(nn.Module):
    def __init__(self, dim, hidden, dropout):
        super().__init__()
        self.lin1, self.lin2 = nn.Linear(dim, hidden), nn.Linear(hidden, dim)
        self.dropout = nn.Dropout(dropout)

    def forward(self, x):
        return self.dropout(self.lin2(F.gelu(self.lin1(x))))


class _Block(nn.Module):
    def __init__(self, dim, heads, hidden, dropout, attn_dropout):
        super().__init__()
        self.attention = _SelfAttention(dim, heads, attn_dropout)
        self.sa_layer_norm = nn.LayerNorm(dim, eps=1e-12)
        self.ffn = _FFN(dim, hidden, dropout)
        self.output_layer_norm = nn.LayerNorm(dim, eps=1e-12)

    def forward(self, x, mask):
        x = self.sa_layer_norm(x + self.attention(x, mask))
        return self.output_layer_norm(x + self.ffn(x))


class _Transformer(nn.Module):
    def __init__(self, n_layers, *a):
        super().__init__()
        self.layer = nn.ModuleList([_Block(*a) for _ in range(n_layers)])

    def forward(self, x, mask):
        for blk in self.layer:
            x = blk(x, mask)
        return x


class DistilBertModel(nn.Module):
    def __init__(self, vocab_size=30522, dim=768, n_layers=6, n_heads=12, hidden_dim=3072, max_pos=512,
                 dropout=0.1, attention_dropout=0.1):
        super().__init__()
        self.embeddings = _Embeddings(vocab_size, dim, max_pos, dropout)
        self.transformer = _Transformer(n_layers, dim, n_heads, hidden_dim, dropout, attention_dropout)

    def forward(self, input_ids, attention_mask=None):
        if attention_mask is None:
            attention_mask = input_ids != 0
        return self.transformer(self.embeddings(input_ids), attention_mask)


class DistilBertForSequenceClassification(nn.Module):
    def __init__(self, num_labels=20, seq_classif_dropout=0.2, **cfg):
        super().__init__()
        dim = cfg.get("dim", 768)
        self.distilbert = DistilBertModel(**cfg)
        self.pre_classifier = nn.Linear(dim, dim)
        self.classifier = nn.Linear(dim, num_labels)
        self.dropout = nn.Dropout(seq_classif_dropout)
        self.apply(self._init)

    @staticmethod
    def _init(m):
        if isinstance(m, (nn.Linear, nn.Embedding)):
            nn.init.normal_(m.weight, std=0.02)
        if isinstance(m, nn.Linear) and m.bias is not None:
            nn.init.zeros_(m.bias)
        if isinstance(m, nn.LayerNorm):
            nn.init.ones_(m.weight); nn.init.zeros_(m.bias)

    def forward(self, input_ids, attention_mask=None):
        h = self.distilbert(input_ids, attention_mask)[:, 0]
        return self.classifier(self.dropout(F.relu(self.pre_classifier(h))))


def distilbert(num_labels=20, tiny=False):
    if tiny:  # CPU-test sized
        return DistilBertForSequenceClassification(num_labels, vocab_size=30522, dim=64, n_layers=2, n_heads=4,
                                                   hidden_dim=128, max_pos=128)
    return DistilBertForSequenceClassification(num_labels)
