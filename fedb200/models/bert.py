"""BERT sequence classifier with the Hugging Face module tree and state_dict key names.

Plain-PyTorch implementation: the CPU path of every entrypoint and the fp32 oracle for the
fused sm_100a trainer (:mod:`fedb200.models.fused_bert`).  Tree and op order follow the
module dump at src/Serverlesscase/serverless_cancer_classification_with_BioBERT.ipynb:526-570
(BertEmbeddings -> 12 x BertLayer -> BertPooler -> Dropout -> classifier); GELU is the erf
form (:553); LayerNorm eps 1e-12; dropout 0.1.  201 tensors for the base shape.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..config import ModelConfig
from .factory import SequenceClassifierOutput


class BertEmbeddings(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.word_embeddings = nn.Embedding(c.vocab_size, c.hidden_size, padding_idx=0)
        self.position_embeddings = nn.Embedding(c.max_position, c.hidden_size)
        self.token_type_embeddings = nn.Embedding(c.type_vocab_size, c.hidden_size)
        self.LayerNorm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)
        self.dropout = nn.Dropout(c.hidden_dropout)

    def forward(self, input_ids, token_type_ids):
        S = input_ids.shape[1]
        pos = torch.arange(S, device=input_ids.device).unsqueeze(0)
        x = self.word_embeddings(input_ids) + self.token_type_embeddings(token_type_ids) + self.position_embeddings(pos)
        return self.dropout(self.LayerNorm(x))


class BertSelfAttention(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.h, self.d = c.num_heads, c.head_dim
        self.query = nn.Linear(c.hidden_size, c.hidden_size)
        self.key = nn.Linear(c.hidden_size, c.hidden_size)
        self.value = nn.Linear(c.hidden_size, c.hidden_size)
        self.dropout = nn.Dropout(c.attention_dropout)

    def forward(self, x, add_mask):
        B, S, H = x.shape
        sp = lambda t: t.view(B, S, self.h, self.d).transpose(1, 2)
        q, k, v = sp(self.query(x)), sp(self.key(x)), sp(self.value(x))
        s = q @ k.transpose(-1, -2) / math.sqrt(self.d) + add_mask
        p = self.dropout(torch.softmax(s, dim=-1))
        return (p @ v).transpose(1, 2).reshape(B, S, H)


class BertSelfOutput(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.dense = nn.Linear(c.hidden_size, c.hidden_size)
        self.LayerNorm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)
        self.dropout = nn.Dropout(c.hidden_dropout)

    def forward(self, h, residual):
        return self.LayerNorm(self.dropout(self.dense(h)) + residual)


class BertAttention(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.self = BertSelfAttention(c)
        self.output = BertSelfOutput(c)

    def forward(self, x, add_mask):
        return self.output(self.self(x, add_mask), x)


class BertIntermediate(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.dense = nn.Linear(c.hidden_size, c.intermediate_size)

    def forward(self, x):
        return F.gelu(self.dense(x))


class BertOutput(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.dense = nn.Linear(c.intermediate_size, c.hidden_size)
        self.LayerNorm = nn.LayerNorm(c.hidden_size, eps=c.layer_norm_eps)
        self.dropout = nn.Dropout(c.hidden_dropout)

    def forward(self, h, residual):
        return self.LayerNorm(self.dropout(self.dense(h)) + residual)


class BertLayer(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.attention = BertAttention(c)
        self.intermediate = BertIntermediate(c)
        self.output = BertOutput(c)

    def forward(self, x, add_mask):
        a = self.attention(x, add_mask)
        return self.output(self.intermediate(a), a)


class BertEncoder(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.layer = nn.ModuleList(BertLayer(c) for _ in range(c.num_layers))

    def forward(self, x, add_mask):
        for l in self.layer:
            x = l(x, add_mask)
        return x


class BertPooler(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.dense = nn.Linear(c.hidden_size, c.hidden_size)

    def forward(self, x):
        return torch.tanh(self.dense(x[:, 0]))


class BertModel(nn.Module):
    def __init__(self, c: ModelConfig):
        super().__init__()
        self.embeddings = BertEmbeddings(c)
        self.encoder = BertEncoder(c)
        self.pooler = BertPooler(c)

    def forward(self, input_ids, token_type_ids, attention_mask):
        add_mask = (1.0 - attention_mask[:, None, None, :].to(torch.float32)) * torch.finfo(torch.float32).min
        x = self.embeddings(input_ids, token_type_ids)
        return self.pooler(self.encoder(x, add_mask))


class BertForSequenceClassification(nn.Module):
    """``AutoModelForSequenceClassification`` for a BERT checkpoint
    (src/Servercase/server_IID_IMDB.py:142-144), built from a config with random init."""

    def __init__(self, config: ModelConfig):
        super().__init__()
        self.config = config
        self.bert = BertModel(config)
        self.dropout = nn.Dropout(config.hidden_dropout)
        self.classifier = nn.Linear(config.hidden_size, config.num_labels)
        self.apply(self._init)

    def _init(self, m):
        std = self.config.initializer_range
        if isinstance(m, nn.Linear):
            nn.init.normal_(m.weight, 0.0, std)
            nn.init.zeros_(m.bias)
        elif isinstance(m, nn.Embedding):
            nn.init.normal_(m.weight, 0.0, std)
            if m.padding_idx is not None:
                with torch.no_grad():
                    m.weight[m.padding_idx].zero_()
        elif isinstance(m, nn.LayerNorm):
            nn.init.ones_(m.weight)
            nn.init.zeros_(m.bias)

    def forward(self, input_ids, token_type_ids: Optional[torch.Tensor] = None,
                attention_mask: Optional[torch.Tensor] = None, labels: Optional[torch.Tensor] = None):
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        if attention_mask is None:
            attention_mask = torch.ones_like(input_ids)
        pooled = self.bert(input_ids, token_type_ids, attention_mask)
        logits = self.classifier(self.dropout(pooled))
        loss = F.cross_entropy(logits, labels) if labels is not None else None
        return SequenceClassifierOutput(loss=loss, logits=logits)
