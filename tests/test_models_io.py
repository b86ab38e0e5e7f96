"""Model factory, HF key compatibility, flat arena, AdamW semantics, client API, checkpoint I/O."""
import math
from collections import OrderedDict

import numpy as np
import pytest
import torch

from fedb200.config import FedConfig, ModelConfig
from fedb200.models import FlatArena, build_model, dir_size_gib, load_pretrained, save_pretrained
from fedb200.models.hf_io import hf_config_dict, load_config, load_state
from fedb200.trainer import FlatAdamW, TorchRunner, hf_adamw_reference_step


def test_bert_base_has_201_tensors_and_reference_param_count():
    c = ModelConfig.preset("biobert", num_labels=40)
    with torch.device("meta"):
        m = build_model(c, "meta")
    sd = m.state_dict()
    assert len(sd) == 201
    assert sum(v.numel() for v in sd.values()) == 108_341_032
    keys = list(sd)
    assert keys[0] == "bert.embeddings.word_embeddings.weight" and keys[-1] == "classifier.bias"
    assert "bert.encoder.layer.11.output.LayerNorm.bias" in sd


def test_albert_has_27_tensors():
    c = ModelConfig.preset("albert-base-v2", num_labels=2)
    with torch.device("meta"):
        m = build_model(c, "meta")
    sd = m.state_dict()
    assert len(sd) == 27 and sum(v.numel() for v in sd.values()) == 11_685_122


@pytest.mark.parametrize("name", ["bert-tiny", "albert-tiny"])
def test_matches_transformers_numerically(name):
    tr = pytest.importorskip("transformers")
    c = ModelConfig.preset(name, num_labels=3)
    m = build_model(c, seed=0).eval()
    kw = {k: v for k, v in hf_config_dict(c).items() if k not in ("model_type", "architectures", "torch_dtype")}
    if c.arch == "bert":
        h = tr.BertForSequenceClassification(tr.BertConfig(**kw))
    else:
        h = tr.AlbertForSequenceClassification(tr.AlbertConfig(**kw))
    assert list(h.state_dict().keys()) == list(m.state_dict().keys())
    h.load_state_dict(m.state_dict())
    h.eval()
    ids = torch.randint(1, 500, (3, 17))
    am = torch.ones(3, 17, dtype=torch.long)
    am[1, 9:] = 0
    y = torch.tensor([0, 1, 2])
    a, b = m(ids, attention_mask=am, labels=y), h(input_ids=ids, attention_mask=am, labels=y)
    torch.testing.assert_close(a.logits, b.logits, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(a.loss, b.loss, rtol=1e-5, atol=1e-6)


def test_flat_arena_views_alias_and_alignment():
    shapes = OrderedDict([("a", (3, 5)), ("b", (7,)), ("c", (2, 2, 2))])
    ar = FlatArena(shapes, key_order=["c", "a", "b"])
    assert all(o % 64 == 0 for o, _ in ar.layout.values())
    ar.views["b"].fill_(2.0)
    o, n = ar.layout["b"]
    assert ar.flat[o:o + n].sum().item() == 14.0
    assert list(ar.state_dict()) == ["c", "a", "b"]
    m = torch.nn.Linear(4, 3)
    arena = FlatArena.adopt(m)
    arena.flat.zero_()
    assert m.weight.abs().sum().item() == 0.0               # parameters are views of the arena


def test_hf_adamw_differs_from_torch_adamw_and_flat_matches_per_tensor():
    torch.manual_seed(0)
    p = torch.randn(1000)
    g = torch.randn(1000) * 1e-3
    p1, m1, v1 = p.clone(), torch.zeros(1000), torch.zeros(1000)
    grads = g.clone()
    flat_p = p.clone()
    opt = FlatAdamW(flat_p, grads, lr=5e-5, use_kernel=False)
    for t in (1, 2, 3):
        grads.copy_(g)
        opt.step(zero_grad=True)
        hf_adamw_reference_step(p1, g, m1, v1, t)
    torch.testing.assert_close(flat_p, p1)
    assert grads.abs().sum().item() == 0.0
    # closed form of step 1: p - lr * sqrt(1-b2)/(1-b1) * (1-b1) g / (sqrt((1-b2) g^2) + eps)
    p2, m2, v2 = p.clone(), torch.zeros(1000), torch.zeros(1000)
    hf_adamw_reference_step(p2, g, m2, v2, 1)
    want = p - 5e-5 * math.sqrt(1 - 0.999) / (1 - 0.9) * (0.1 * g) / ((0.001 * g * g).sqrt() + 1e-6)
    torch.testing.assert_close(p2, want)
    q = torch.nn.Parameter(p.clone())
    q.grad = g.clone()
    torch.optim.AdamW([q], lr=5e-5, eps=1e-6, weight_decay=0.0).step()
    assert (q.detach() - p2).abs().max().item() > 1e-7      # eps placement differs


def test_client_api_conventions_and_roundtrip():
    from fedb200.client import FedClient
    from fedb200.data import BatchLoader, make_synthetic
    fed = FedConfig()
    fed.model = ModelConfig.preset("bert-tiny", num_labels=2)
    r = TorchRunner(fed.model, fed, seed=0)
    ds = make_synthetic("imdb", "train", 512, 48, rows=70)
    cl = FedClient(r, BatchLoader(ds, 32, shuffle=True), BatchLoader(ds, 32))
    params = cl.get_parameters({})
    assert len(params) == len(r.state_dict()) and all(p.dtype == np.float32 for p in params)
    new, n, metrics = cl.fit(params, {})
    assert n == 3 and metrics == {}                           # weight = number of BATCHES (70 samples / 32)
    assert any(not np.allclose(a, b) for a, b in zip(params, new))
    loss, n_eval, m = cl.evaluate(new, {})
    assert n_eval == 3 and set(m) == {"accuracy", "loss"} and 0.0 <= m["accuracy"] <= 1.0
    cl.set_parameters(params)
    for a, b in zip(params, cl.get_parameters({})):
        np.testing.assert_array_equal(a, b)
    with pytest.raises(ValueError):
        cl.set_parameters(params[:-1])


def test_eval_loss_uses_reference_mixed_normalisation():
    from fedb200.data import BatchLoader, make_synthetic
    fed = FedConfig()
    fed.model = ModelConfig.preset("bert-tiny", num_labels=2, hidden_dropout=0.0, attention_dropout=0.0)
    r = TorchRunner(fed.model, fed, seed=0)
    ds = make_synthetic("imdb", "train", 512, 32, rows=40)
    loss, acc = r.evaluate(BatchLoader(ds, 32))
    r.module.eval()
    with torch.no_grad():
        per_batch = [float(r.module(**b).loss) for b in BatchLoader(ds, 32)]
    assert abs(loss - sum(per_batch) / 40) < 1e-6            # sum of batch means / len(dataset)


def test_checkpoint_save_load_equality_and_hf_layout(tmp_path):
    c = ModelConfig.preset("bert-tiny", num_labels=5)
    m = build_model(c, seed=1)
    size = save_pretrained(m.state_dict(), c, str(tmp_path / "ck"), {"round": 3})
    assert size > 0 and abs(size - dir_size_gib(str(tmp_path / "ck"))) < 1e-12
    assert (tmp_path / "ck" / "config.json").exists() and (tmp_path / "ck" / "model.safetensors").exists()
    sd = load_state(str(tmp_path / "ck"))
    assert set(sd) == set(m.state_dict())
    m2 = build_model(load_config(str(tmp_path / "ck")), seed=2)
    meta = load_pretrained(m2, str(tmp_path / "ck"))
    assert meta["round"] == 3
    for k, v in m.state_dict().items():
        torch.testing.assert_close(m2.state_dict()[k], v)
    tr = pytest.importorskip("transformers")
    h = tr.AutoModelForSequenceClassification.from_pretrained(str(tmp_path / "ck"))
    torch.testing.assert_close(h.state_dict()["classifier.weight"], m.state_dict()["classifier.weight"])
    m3 = build_model(ModelConfig.preset("bert-tiny", num_labels=7), seed=3)
    load_pretrained(m3, str(tmp_path / "ck"), ignore_mismatched_sizes=True)      # head re-initialised, body loaded
    torch.testing.assert_close(m3.state_dict()["bert.pooler.dense.weight"], m.state_dict()["bert.pooler.dense.weight"])
    assert m3.state_dict()["classifier.weight"].shape[0] == 7


# ---- fine-tuning from a pretrained directory: from_pretrained semantics + WordPiece -------------------------------
_VOCAB = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]", "the", "patient", "##s", "pre", "##sent", "##ed", "with", "a",
          "chest", "pain", ",", ".", "Pat", "##ient", "cafe", "un", "##believ", "##able", "!", "(", ")", "12", "##3", "mg", "/", "d"]


def test_wordpiece_matches_transformers_bert_tokenizer(tmp_path):
    from fedb200.data.tokenizer import HashTokenizer, WordPieceTokenizer, load_tokenizer
    (tmp_path / "vocab.txt").write_text("\n".join(_VOCAB) + "\n", encoding="utf-8")
    texts = ["The patients presented with a chest pain, unbelievable!", "Patient (123 mg/d) café. xyzzy the",
             "  the\tpatient\n\npresented  ", "Patients unbelievable123", ""]
    tr = pytest.importorskip("transformers")
    for lower in (False, True):
        (tmp_path / "tokenizer_config.json").write_text('{"do_lower_case": %s}' % ("true" if lower else "false"))
        mine = WordPieceTokenizer.from_pretrained(str(tmp_path))
        assert mine.lowercase is lower and mine.cls_id == 2 and mine.sep_id == 3 and mine.unk_id == 1
        hf = tr.BertTokenizer(vocab={w: i for i, w in enumerate(_VOCAB)}, do_lower_case=lower)
        for t in texts:
            assert mine.encode(t) == hf(t)["input_ids"], (lower, t)
    # greedy longest match, '##' continuation, whole-word [UNK]
    cased = WordPieceTokenizer({w: i for i, w in enumerate(_VOCAB)}, lowercase=False)
    assert cased.encode("unbelievable patients") == [2, 20, 21, 22, 6, 7, 3]
    assert cased.encode("unbelievablez") == [2, 1, 3]
    assert len(cased.encode("the " * 600)) == 512                                    # truncation=True
    # cased / uncased inferred from the vocabulary when there is no tokenizer_config.json
    (tmp_path / "tokenizer_config.json").unlink()
    assert WordPieceTokenizer.from_pretrained(str(tmp_path)).lowercase is False       # "Pat" is in the vocabulary
    assert isinstance(load_tokenizer(None, 28996), HashTokenizer)
    assert isinstance(load_tokenizer(str(tmp_path), 28996), WordPieceTokenizer)


def test_from_pretrained_semantics_missing_unexpected_renamed_keys(tmp_path):
    """A pre-training checkpoint (BioBERT-style: ``cls.predictions.*`` head, TF LayerNorm names, no classifier) loads the
    way ``AutoModelForSequenceClassification.from_pretrained`` loads it: body taken, extra head ignored, classifier fresh."""
    c = ModelConfig.preset("bert-tiny", num_labels=2)
    src = build_model(c, seed=11)
    sd = OrderedDict((k, v.clone()) for k, v in src.state_dict().items() if not k.startswith("classifier."))
    sd["cls.predictions.bias"] = torch.zeros(c.vocab_size)
    for k in [k for k in sd if k.endswith("LayerNorm.weight")]:
        sd[k[: -len("weight")] + "gamma"] = sd.pop(k)
    for k in [k for k in sd if k.endswith("LayerNorm.bias")]:
        sd[k[: -len("bias")] + "beta"] = sd.pop(k)
    save_pretrained(sd, c, str(tmp_path / "pre"))
    dst = build_model(ModelConfig.preset("bert-tiny", num_labels=40), seed=12)
    fresh_head = dst.state_dict()["classifier.weight"].clone()
    rep = load_pretrained(dst, str(tmp_path / "pre"))["load_report"]
    assert rep["unexpected"] == ["cls.predictions.bias"] and sorted(rep["missing"]) == ["classifier.bias", "classifier.weight"]
    for k, v in src.state_dict().items():
        if not k.startswith("classifier."):
            torch.testing.assert_close(dst.state_dict()[k], v)
    torch.testing.assert_close(dst.state_dict()["classifier.weight"], fresh_head)
    with pytest.raises(KeyError):
        load_pretrained(dst, str(tmp_path / "pre"), strict=True)
    # a classifier of another width is an error unless ignore_mismatched_sizes
    save_pretrained(src.state_dict(), c, str(tmp_path / "two"))
    with pytest.raises(RuntimeError, match="size mismatch"):
        load_pretrained(dst, str(tmp_path / "two"))
    assert sorted(load_pretrained(dst, str(tmp_path / "two"), ignore_mismatched_sizes=True)["load_report"]["mismatched"]) == \
        ["classifier.bias", "classifier.weight"]


def test_init_checkpoint_flag_fine_tunes_loaded_weights_on_csv_text(tmp_path):
    """``serverless_IID.py --source csv --init-checkpoint DIR``: architecture from DIR/config.json, weights from DIR,
    word pieces from DIR/vocab.txt; the run starts from the loaded weights (save_pretrained -> --init-checkpoint cycle)."""
    import csv
    from fedb200.config import build_parser, config_from_args
    from fedb200.runtime import FedRuntime
    words = ["w%d" % i for i in range(200)]
    vocab = ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + words + ["##x"]
    c = ModelConfig.preset("bert-tiny", num_labels=2, vocab_size=len(vocab) + 10)
    pre = build_model(c, seed=21)
    ck = tmp_path / "pretrained"
    save_pretrained(pre.state_dict(), c, str(ck))
    (ck / "vocab.txt").write_text("\n".join(vocab) + "\n")
    data = tmp_path / "Dataset"
    data.mkdir()
    rng = np.random.default_rng(0)
    for fn, n in (("train_file_mt.csv", 120), ("test_file_mt.csv", 60)):
        with open(data / fn, "w", newline="") as fh:
            w = csv.writer(fh)
            w.writerow(["", "description", "medical_specialty"])
            for i in range(n):
                y = int(rng.integers(0, 40))
                w.writerow([i, " ".join(words[(y * 5 + int(j)) % 200] for j in rng.integers(0, 5, size=12)), y])
    args = build_parser().parse_args(["--source", "csv", "--dataset", "medical_transcriptions", "--data-dir", str(data),
                                      "--init-checkpoint", str(ck), "--ignore-mismatched-sizes", "--clients", "2",
                                      "--rounds", "1", "--train-per-client", "32", "--test-per-client", "16",
                                      "--device", "cpu", "--backend", "torch", "--lr", "1e-3"])
    cfg = config_from_args(args)
    assert cfg.model.hidden_size == 64 and cfg.model.num_labels == 40 and cfg.model.vocab_size == len(vocab) + 10
    rt = FedRuntime(cfg, echo=False)
    sd0 = rt.runner.state_dict()
    torch.testing.assert_close(sd0["bert.encoder.layer.0.attention.self.query.weight"].cpu(),
                               pre.state_dict()["bert.encoder.layer.0.attention.self.query.weight"])
    assert sd0["classifier.weight"].shape[0] == 40                       # mismatched head re-initialised
    assert int(rt.corpus.train.tokens.max()) < len(vocab) and int(rt.corpus.train.sequence(0)[0]) == 2    # real vocab ids
    out = rt.run()
    rt.close()
    assert len(out["global_accuracies"]) == 1
    moved = (rt.runner.state_dict()["bert.encoder.layer.0.attention.self.query.weight"].cpu()
             - pre.state_dict()["bert.encoder.layer.0.attention.self.query.weight"]).abs().max()
    assert 0 < float(moved) < 0.1                                        # fine-tuned FROM the loaded weights
