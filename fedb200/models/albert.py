"""ALBERT-base-v2 sequence classifier (HF key names; 27 tensors, 11.7 M parameters).

Used by the reference's IMDB / cancer / covid scripts (``CHECKPOINT = "albert-base-v2"``,
src/Serverlesscase/serverless_IID_IMDB.py:31).  One shared transformer layer applied
``num_layers`` times, factorised 128-d embeddings, ``gelu_new`` (tanh form) activation,
post-LN residual blocks; encoder dropouts are 0 in albert-base-v2, classifier dropout 0.1.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..config import ModelConfig
from .factory import SequenceClassifierOutput


class AlbertEmbeddings(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        E = c.embedding_size
        self.word_embeddings = nn.Embedding(c.vocab_size, E, padding_idx=0)
        self.position_embeddings = nn.Embedding(c.max_position, E)
        self.token_type_embeddings = nn.Embedding(c.type_vocab_size, E)
        self.LayerNorm = nn.LayerNorm(E, eps=c.layer_norm_eps)

    def forward(self, input_ids, token_type_ids):
        pos = torch.arange(input_ids.shape[1], device=input_ids.device).unsqueeze(0)
        return self.LayerNorm(self.word_embeddings(input_ids) + self.token_type_embeddings(token_type_ids)
                              + self.position_embeddings(pos))


class AlbertAttention(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.h, self.d = c.num_heads, c.head_dim
        H = c.hidden_size
        self.query, self.key, self.value = nn.Linear(H, H), nn.Linear(H, H), nn.Linear(H, H)
        self.dense = nn.Linear(H, H)
        self.LayerNorm = nn.LayerNorm(H, eps=c.layer_norm_eps)

    def forward(self, x, add_mask):
        B, S, H = x.shape
        sp = lambda t: t.view(B, S, self.h, self.d).transpose(1, 2)
        q, k, v = sp(self.query(x)), sp(self.key(x)), sp(self.value(x))
        p = torch.softmax(q @ k.transpose(-1, -2) / math.sqrt(self.d) + add_mask, dim=-1)
        ctx = (p @ v).transpose(1, 2).reshape(B, S, H)
        return self.LayerNorm(x + self.dense(ctx))


class AlbertLayer(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.full_layer_layer_norm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)
        self.attention = AlbertAttention(c)
        self.ffn = nn.Linear(c.hidden_size, c.intermediate_size)
        self.ffn_output = nn.Linear(c.intermediate_size, c.hidden_size)

    def forward(self, x, add_mask):
        a = self.attention(x, add_mask)
        f = self.ffn_output(F.gelu(self.ffn(a), approximate="tanh"))
        return self.full_layer_layer_norm(f + a)


class AlbertLayerGroup(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.albert_layers = nn.ModuleList([AlbertLayer(c)])

    def forward(self, x, add_mask):
        for l in self.albert_layers:
            x = l(x, add_mask)
        return x


class AlbertTransformer(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.n = c.num_layers
        self.embedding_hidden_mapping_in = nn.Linear(c.embedding_size, c.hidden_size)
        self.albert_layer_groups = nn.ModuleList([AlbertLayerGroup(c)])

    def forward(self, x, add_mask):
        x = self.embedding_hidden_mapping_in(x)
        for _ in range(self.n):
            x = self.albert_layer_groups[0](x, add_mask)
        return x


class AlbertModel(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.embeddings = AlbertEmbeddings(c)
        self.encoder = AlbertTransformer(c)
        self.pooler = nn.Linear(c.hidden_size, c.hidden_size)

    def forward(self, input_ids, token_type_ids, attention_mask):
        add_mask = (1.0 - attention_mask[:, None, None, :].to(torch.float32)) * torch.finfo(torch.float32).min
        x = self.encoder(self.embeddings(input_ids, token_type_ids), add_mask)
        return torch.tanh(self.pooler(x[:, 0]))


class AlbertForSequenceClassification(nn.Module):
    def __init__(self, config: ModelConfig):
        super().__init__()
        self.config = config
        self.albert = AlbertModel(config)
        self.dropout = nn.Dropout(0.1 if config.hidden_dropout > 0 else 0.0)
        self.classifier = nn.Linear(config.hidden_size, config.num_labels)
        self.apply(self._init)

    def _init(self, m):
        std = self.config.initializer_range
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, 0.0, std)
            nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, 0.0, std)
        elif isinstance(m, nn.LayerNorm):
            nn.init.ones_(m.weight)
            nn.init.zeros_(m.bias)

    def forward(self, input_ids, token_type_ids: Optional[torch.Tensor] = None,
                attention_mask: Optional[torch.Tensor] = None, labels: Optional[torch.Tensor] = None):
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        if attention_mask is None:
            attention_mask = torch.ones_like(input_ids)
        pooled = self.albert(input_ids, token_type_ids, attention_mask)
        logits = self.classifier(self.dropout(pooled))
        loss = F.cross_entropy(logits, labels) if labels is not None else None
        return SequenceClassifierOutput(loss=loss, logits=logits)
