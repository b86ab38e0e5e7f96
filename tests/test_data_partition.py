"""Data pipeline + partitioners (C1-C3, C27): index math of the reference's splits, Dirichlet skew, collation."""
import numpy as np
import torch

from fedb200.config import DataConfig, ModelConfig
from fedb200.data import (BatchLoader, HashTokenizer, Partitioner, TokenDataset, collate_padded, dirichlet_indices,
                          iid_indices, make_synthetic, shard_indices)
from fedb200.data.build import build_corpus, client_loaders, global_test_loader
from fedb200.data.dataset import collate_packed


def test_iid_sample_is_distinct_and_sized():
    rng = np.random.default_rng(0)
    idx = iid_indices(25000, 100, rng)
    assert len(idx) == 100 == len(set(idx.tolist())) and idx.max() < 25000
    assert len(iid_indices(50, 100, rng)) == 50           # random.sample cannot exceed the population


def test_shard_index_math_matches_reference_strides():
    tr, te = shard_indices(3, 300, 240, "tail")           # serverless_NonIID_IMDB.py:59-60
    assert tr[0] == 900 and tr[-1] == 1139 and len(tr) == 240
    assert te[0] == 1140 and te[-1] == 1199 and len(te) == 60
    tr, te = shard_indices(2, 500, 400, "fixed", 400)     # Serverless_NonIID_Medical_transcriptions.py:55-56
    assert tr[0] == 1000 and tr[-1] == 1399
    assert te[0] == 0 and te[-1] == 399


def test_imdb_shards_are_single_class_but_medical_are_not():
    imdb = make_synthetic("imdb", "train", 2000, 128, rows=6000, shuffled=False)
    tr, _ = shard_indices(0, 300, 240)
    assert len(set(imdb.labels[tr].tolist())) == 1        # label-sorted split => pathological skew
    med = make_synthetic("medical_transcriptions", "train", 2000, 128, rows=6000, shuffled=False)
    tr, _ = shard_indices(0, 500, 400)
    assert len(set(med.labels[tr].tolist())) > 20


def test_dirichlet_split_is_a_partition_and_skewed():
    rng = np.random.default_rng(1)
    labels = rng.integers(0, 4, size=4000)
    parts = dirichlet_indices(labels, 8, 0.1, np.random.default_rng(2))
    allidx = np.concatenate(parts)
    assert len(allidx) == 4000 and len(set(allidx.tolist())) == 4000
    fracs = [np.bincount(labels[p], minlength=4).max() / len(p) for p in parts]
    assert np.mean(fracs) > 0.6                            # strong label skew at alpha = 0.1
    even = dirichlet_indices(labels, 8, 1000.0, np.random.default_rng(2))
    assert np.mean([np.bincount(labels[p], minlength=4).max() / len(p) for p in even]) < 0.4


def test_partitioner_resamples_iid_every_round_and_is_deterministic():
    p = Partitioner("iid", 25000, 25000, 100, 100, seed=7)
    a, _ = p.indices(3, 0)
    b, _ = p.indices(3, 1)
    c, _ = p.indices(3, 0)
    assert not np.array_equal(a, b) and np.array_equal(a, c)


def test_padded_and_packed_collation_agree():
    ds = TokenDataset.from_sequences([[101, 5, 6, 102], [101, 7, 102], [101, 8, 9, 10, 11, 102]], [0, 1, 0], 2)
    idx = np.array([0, 1, 2])
    pad = collate_padded(ds, idx)
    assert pad["input_ids"].shape == (3, 6) and pad["attention_mask"].sum().item() == 13
    assert pad["input_ids"][1].tolist() == [101, 7, 102, 0, 0, 0]
    pk = collate_packed(ds, idx)
    assert pk["cu_seqlens"].tolist() == [0, 4, 7, 13] and pk["num_tokens"] == 13 and pk["max_len"] == 6
    assert pk["ids"].tolist() == [101, 5, 6, 102, 101, 7, 102, 101, 8, 9, 10, 11, 102]
    assert pk["pos"].tolist() == [0, 1, 2, 3, 0, 1, 2, 0, 1, 2, 3, 4, 5]
    flat = pad["input_ids"][pad["attention_mask"].bool()]
    assert flat.tolist() == pk["ids"].tolist()


def test_loader_len_is_number_of_batches_and_shuffle_only_train():
    ds = make_synthetic("imdb", "train", 2000, 64, rows=100)
    tr = BatchLoader(ds, 32, shuffle=True, rng=np.random.default_rng(0))
    te = BatchLoader(ds, 32, shuffle=False)
    assert len(tr) == 4 and len(te) == 4 and tr.num_samples == 100       # 32,32,32,4
    sizes = [b["labels"].numel() for b in te]
    assert sizes == [32, 32, 32, 4]
    first = next(iter(te))["labels"]
    assert torch.equal(first, torch.from_numpy(ds.labels[:32]))


def test_corpus_and_client_loaders_all_split_kinds():
    m = ModelConfig.preset("bert-tiny")
    for split in ("iid", "shard", "dirichlet"):
        d = DataConfig(split=split, population=3000, train_per_client=48, test_per_client=16, max_seq_len=64)
        c = build_corpus(d, m, 4, seed=1)
        tr, te = client_loaders(c, d, 2, 0)
        assert tr.num_samples > 0 and te.num_samples > 0
        b = next(iter(tr))
        assert b["input_ids"].shape[1] <= 64 and set(b) == {"input_ids", "token_type_ids", "attention_mask", "labels"}
        g = global_test_loader(c, d, 0)
        assert g.num_samples == 16


def test_hash_tokenizer_structure_and_truncation():
    tok = HashTokenizer(vocab_size=28996, max_length=16)
    ids = tok.encode("The patient presented with acute abdominal pain, nausea and vomiting for three days running.")
    assert ids[0] == 101 and ids[-1] == 102 and len(ids) == 16
    assert all(1000 <= t < 28996 for t in ids[1:-1])
    assert tok.encode("pain") == tok.encode("pain")


def test_bundled_csv_loader(tmp_path):
    from fedb200.data import load_csv_dataset
    p = tmp_path / "train_file_mt.csv"
    p.write_text("Unnamed: 0,description,medical_specialty\n0,Chest pain and dyspnea,3\n1,Knee arthroscopy,27\n")
    (tmp_path / "test_file_mt.csv").write_text("Unnamed: 0,description,medical_specialty\n0,Follow up visit,5\n")
    ds = load_csv_dataset("medical_transcriptions", str(tmp_path), "train", 28996, 64)
    assert len(ds) == 2 and ds.labels.tolist() == [3, 27] and ds.num_labels == 40
