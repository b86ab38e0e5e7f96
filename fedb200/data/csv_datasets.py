"""Loaders for the CSV files the reference bundles under ``Dataset/`` (C27).

No reference script reads these files (they all pull HF hub ids), but they are the only
real data available offline: ``train_file_mt.csv``/``test_file_mt.csv`` (columns
``description, medical_specialty``; 40 integer classes) and the self-driving sentiment
CSVs (``Text, Sentiment``).  Column conventions follow
src/Servercase/server_iid_medical_transcirptions.py:56,68.
"""
from __future__ import annotations

import csv
import os
from typing import Optional

from .dataset import TokenDataset
from .tokenizer import load_tokenizer

CSV_SPECS = {
    "medical_transcriptions": dict(train="train_file_mt.csv", test="test_file_mt.csv",
                                   text="description", label="medical_specialty", num_labels=40),
    "self_driving": dict(train="sentiment_analysis_self_driving_vehicles.csv",
                         test="sentiment_analysis_self_driving_vehicles.csv",
                         text="Text", label="Sentiment", num_labels=3),
    "self_driving_ctgan": dict(train="Augmeted_datasets/CTGAN_self_driving_vehicles.csv",
                               test="sentiment_analysis_self_driving_vehicles.csv",
                               text="Text", label="Sentiment", num_labels=3),
    "self_driving_copula": dict(train="Augmeted_datasets/output_Gaussiancopula_self_driving.csv",
                                test="sentiment_analysis_self_driving_vehicles.csv",
                                text="Text", label="Sentiment", num_labels=3),
}


def load_csv_dataset(name: str, data_dir: str, split: str = "train", vocab_size: int = 28996,
                     max_seq_len: int = 512, rows: Optional[int] = None, vocab_path: Optional[str] = None) -> TokenDataset:
    """``vocab_path``: checkpoint directory / ``vocab.txt`` -> real WordPiece ids (``AutoTokenizer.from_pretrained``,
    src/Servercase/server_IID_IMDB.py:72-76); without one the hash tokenizer stands in."""
    spec = CSV_SPECS[name]
    path = os.path.join(data_dir, spec[split if split in ("train", "test") else "train"])
    tok = load_tokenizer(vocab_path, vocab_size, max_seq_len)
    if getattr(tok, "vocab_size", vocab_size) > vocab_size:
        raise ValueError(f"tokenizer vocabulary ({tok.vocab_size}) exceeds the model's embedding table ({vocab_size})")
    seqs, labels, label_map = [], [], {}
    with open(path, newline="", encoding="utf-8", errors="replace") as fh:
        for i, row in enumerate(csv.DictReader(fh)):
            if rows is not None and i >= rows:
                break
            text, lab = row.get(spec["text"]), row.get(spec["label"])
            if text is None or lab is None:
                continue
            try:
                y = int(float(lab))
            except ValueError:
                y = label_map.setdefault(lab.strip().lower(), len(label_map))
            seqs.append(tok.encode(text))
            labels.append(y)
    return TokenDataset.from_sequences(seqs, labels, spec["num_labels"])
