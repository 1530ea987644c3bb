"""SURVEY 8f rows N1 (text -> ids), N2 (embedding cache), N4 (optimizer tables, checkpoints): oracle pinned to the
reference's goldens, native host tokenizer vs oracle (CPU - the tokenizer is host code of the C-ABI library), and the
GPU legs (fused clip + Adam/AdamW vs torch.optim, pinned id ring, search service vs the oracle)."""
import json
import os
import random

import numpy as np
import pytest
import torch

from conftest import GOLDEN, rel_err
from oracle import ref_pipeline as rpl
from oracle.ref_import import load_reference, reference_available

import two_towers_b200 as tt
from two_towers_b200 import optim as topt

G = json.load(open(os.path.join(GOLDEN, "pipeline.json")))
INDEX = {}
for _i, _w in enumerate(G["words"]):
    INDEX.setdefault(_w, _i + 1)


# ------------------------------------------------------------------------------------------ oracle pinned
@pytest.mark.parametrize("L", G["max_lengths"])
def test_oracle_matches_reference_goldens(L):
    for text, s, e, m in zip(G["texts"], G["simple"][str(L)], G["enhanced"][str(L)], G["margin"][str(L)]):
        assert rpl.simple_ids(text, INDEX, L) == s == e
        assert rpl.margin_ids(text, INDEX, L) == m


@pytest.mark.skipif(not reference_available(), reason="reference tree not present (GPU box)")
def test_oracle_matches_live_reference():
    simple, margin, enhanced = load_reference()

    class KV:
        vector_size = 2

        def __getitem__(self, w):
            return np.array([INDEX[w], 0], dtype=np.float32)

    rng = random.Random(5)
    pool = G["words"] + ["zz", "q1", "12", "kg", "mg", "Part", "OF", "operates", "regulates", "3.5"]
    for _ in range(200):
        text = " ".join(rng.choice(pool) for _ in range(rng.randint(0, 40)))
        L = rng.choice([3, 10, 30])
        ref_s = simple.SimpleDataset([], [], KV(), max_length=L).text_to_embedding(text)[:, 0].long().tolist()
        ref_m = margin.SimpleDataset.text_to_embedding(text, KV(), L)[:, 0].long().tolist()
        assert rpl.simple_ids(text, INDEX, L) == ref_s
        assert rpl.margin_ids(text, INDEX, L) == ref_m


# ------------------------------------------------------------------------------------------ native tokenizer (host code)
@pytest.fixture(scope="module")
def vocab():
    return tt.Vocabulary(G["words"], pad_id=0, first_id=1)


@pytest.mark.parametrize("L", G["max_lengths"])
@pytest.mark.parametrize("dtype", [torch.int64, torch.int32])
def test_native_tokenizer_matches_goldens(vocab, L, dtype):
    out = torch.full((len(G["texts"]), L), -7, dtype=dtype)
    lens = torch.zeros(len(G["texts"]), dtype=torch.int32)
    vocab.encode_batch(G["texts"], L, out=out, lengths=lens, threads=3)
    assert out.tolist() == G["simple"][str(L)]
    assert lens.tolist() == [sum(1 for v in row if v != 0) for row in G["simple"][str(L)]]


def test_margin_encoding_matches_goldens(vocab):
    for L in G["max_lengths"]:
        assert vocab.encode_margin_batch(G["texts"], L).tolist() == G["margin"][str(L)]


def test_native_tokenizer_random_texts_vs_oracle(vocab):
    rng = random.Random(11)
    spaces = [" ", "  ", "\t", "\n", "\r\n", "\u00a0", "\u2003", "\u3000", "\x1c", "\u2028", "\u0085 "]
    pool = G["words"] + ["OOV%d" % i for i in range(30)] + ["The", "PASTA", "\u00c7a", "x\u200by"]
    texts = []
    for _ in range(3000):
        n = rng.randint(0, 50)
        t = rng.choice(["", " ", "\n"]) + "".join(rng.choice(pool) + rng.choice(spaces) for _ in range(n))
        texts.append(t if rng.random() < 0.7 else t.rstrip())
    for L, threads in ((7, 1), (32, 0), (128, 5)):
        got = vocab.encode_batch(texts, L, threads=threads)
        want = [rpl.simple_ids(t, INDEX, L) for t in texts]
        assert got.tolist() == want
        # mode 1 (additive): drop unknown words first, then truncate
        got1 = vocab.encode_batch(texts, L, mode=1)
        want1 = []
        for t in texts:
            ids = [INDEX[w] for w in t.lower().split() if w in INDEX][:L]
            want1.append(ids + [0] * (L - len(ids)))
        assert got1.tolist() == want1


def test_table_gather_equals_reference_embedding(vocab):
    """table[ids] == text_to_embedding(text): rows of the word vectors, zero rows as padding."""
    E = 6
    g = torch.Generator().manual_seed(0)
    table = torch.randn(vocab.table_rows(), E, generator=g)
    table[0] = 0
    L = 10
    ids = vocab.encode_batch(G["texts"], L)
    emb = table[ids]
    for b, row in enumerate(G["simple"][str(L)]):
        want = torch.stack([table[i] if i else torch.zeros(E) for i in row])
        assert torch.equal(emb[b], want)


def test_vocabulary_argument_errors(vocab):
    with pytest.raises(ValueError):
        tt.Vocabulary(["a", "b"], pad_id=1, first_id=1)
    with pytest.raises(ValueError):
        vocab.encode_batch(["a"], 4, out=torch.zeros(2, 4, dtype=torch.int64))
    with pytest.raises(ValueError):
        vocab.encode_batch(["a"], 4, out=torch.zeros(1, 4, dtype=torch.float32))
    assert vocab.encode_batch([], 4).shape == (0, 4)
    assert len(vocab) == len(INDEX) and "pasta" in vocab and "zzz" not in vocab


def test_from_word2vec_layout():
    class KV:
        index_to_key = ["x", "y", "z"]
        vectors = np.arange(12, dtype=np.float32).reshape(3, 4)

    v, table = tt.Vocabulary.from_word2vec(KV())
    assert table.shape == (4, 4) and torch.equal(table[0], torch.zeros(4)) and table[2, 0] == 4
    assert v.encode_batch(["z x q y"], 5).tolist() == [[3, 1, 2, 0, 0]]


# ------------------------------------------------------------------------------------------ optimizer tables, checkpoints
def test_opt_tables_cover_every_element_once():
    entries = [(1000, 2000, 3000, 4000, 5), (10000, 20000, 30000, 40000, 65536), (7, 8, 9, 10, 200001), (1, 2, 3, 4, 0)]
    tb, cb, n = topt.build_tables(entries)
    import ctypes as C
    assert tb.numel() == len(entries) * C.sizeof(topt.OptTensor) == len(entries) * 40
    assert cb.numel() == n * C.sizeof(topt.OptChunk) == n * 16
    ck = np.frombuffer(cb.numpy().tobytes(), dtype=[("t", "<i4"), ("r", "<i4"), ("o", "<i8")])
    cover = {}
    for t, _, o in ck:
        numel = entries[t][4]
        assert o % 4 == 0 and 0 <= o < numel
        cover[t] = cover.get(t, 0) + min(topt.OPT_CHUNK, numel - o)
    assert cover == {0: 5, 1: 65536, 2: 200001}
    tn = np.frombuffer(tb.numpy().tobytes(), dtype=[("p", "<u8"), ("g", "<u8"), ("m", "<u8"), ("v", "<u8"), ("n", "<i8")])
    assert tn["p"].tolist() == [1000, 10000, 7, 1] and tn["n"].tolist() == [5, 65536, 200001, 0]


def test_checkpoint_formats_roundtrip(tmp_path):
    model = tt.EnhancedTwoTowerModel(20, 4)
    ref_opt = torch.optim.AdamW(model.parameters(), lr=1e-4, weight_decay=0.01)
    p = str(tmp_path / "best_model.pt")
    tt.save_checkpoint(p, model, ref_opt, epoch=3, val_loss=0.25)       # simple_two_tower.py:263-268
    ck = torch.load(p)
    assert set(ck) == {"epoch", "model_state_dict", "optimizer_state_dict", "val_loss"}
    m2 = tt.EnhancedTwoTowerModel(20, 4)
    o2 = tt.FusedAdamW(m2.parameters(), lr=1e-4, weight_decay=0.01, max_grad_norm=1.0)
    got = tt.load_checkpoint(p, m2, o2)
    assert got["epoch"] == 3 and got["val_loss"] == 0.25
    for (k, a), (_, b) in zip(model.state_dict().items(), m2.state_dict().items()):
        assert torch.equal(a, b), k
    assert o2.param_groups[0]["lr"] == 1e-4 and o2.param_groups[0]["weight_decay"] == 0.01
    # and the other way: our optimizer's state_dict loads into the reference's optimizer class
    ref_opt.load_state_dict(o2.state_dict())
    # bare state_dict form (train_enhanced.py:76)
    tt.save_checkpoint(p, model)
    assert "model_state_dict" not in torch.load(p)
    assert tt.load_checkpoint(p, m2) == {}


def test_optimizer_refuses_cpu_parameters():
    w = torch.nn.Parameter(torch.ones(4))
    w.grad = torch.ones(4)
    with pytest.raises(RuntimeError, match="CUDA"):
        tt.FusedAdam([w]).step()


# ------------------------------------------------------------------------------------------ GPU legs
def _rand_params(dev, seed, shapes):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter((torch.randn(*s, generator=g) * 0.3).to(dev)) for s in shapes]


SHAPES = [(768, 300), (768, 256), (768,), (3,), (1,), (70001,), (256, 513)]


@pytest.mark.gpu
@pytest.mark.parametrize("kind,wd,clip", [("adam", 0.0, None), ("adam", 0.0, 1.0), ("adam", 0.05, 0.5),
                                          ("adamw", 0.01, 1.0), ("adamw", 0.01, None), ("adamw", 0.0, 1e6)])
def test_fused_adam_matches_torch(kind, wd, clip):
    dev = torch.device("cuda:0")
    ours_p, ref_p = _rand_params(dev, 1, SHAPES), _rand_params(dev, 1, SHAPES)
    if kind == "adam":
        ours = tt.FusedAdam(ours_p, lr=1e-3, weight_decay=wd, max_grad_norm=clip)
        ref = torch.optim.Adam(ref_p, lr=1e-3, weight_decay=wd)
    else:
        ours = tt.FusedAdamW(ours_p, lr=1e-4, weight_decay=wd, max_grad_norm=clip)
        ref = torch.optim.AdamW(ref_p, lr=1e-4, weight_decay=wd)
    g = torch.Generator().manual_seed(2)
    for step in range(6):
        scale = [3.0, 0.01, 1.0, 40.0, 1e-3, 2.0][step]
        for a, b in zip(ours_p, ref_p):
            gr = (torch.randn(a.shape, generator=g) * scale).to(dev)
            a.grad, b.grad = gr.clone(), gr.clone()
        if clip is not None:
            want_norm = torch.nn.utils.clip_grad_norm_(ref_p, max_norm=clip)
        ref.step()
        ours.step()
        if clip is not None:
            assert abs(ours.last_grad_norm.item() - want_norm.item()) <= 2e-6 * want_norm.item()
            for a, b in zip(ours_p, ref_p):                       # gradients are left clipped in place
                assert rel_err(a.grad, b.grad) < 2e-6
        for a, b in zip(ours_p, ref_p):
            assert torch.allclose(a, b, rtol=2e-6, atol=2e-7), (step, a.shape, (a - b).abs().max().item())
    for a, b in zip(ours_p, ref_p):
        assert rel_err(ours.state[a]["exp_avg"], ref.state[b]["exp_avg"]) < 1e-6
        assert rel_err(ours.state[a]["exp_avg_sq"], ref.state[b]["exp_avg_sq"]) < 1e-6
        assert float(ours.state[a]["step"]) == 6.0


@pytest.mark.gpu
def test_fused_adam_param_groups_share_one_clip_norm_and_resume():
    dev = torch.device("cuda:0")
    ours_p, ref_p = _rand_params(dev, 3, SHAPES), _rand_params(dev, 3, SHAPES)
    ours = tt.FusedAdamW([{"params": ours_p[:3], "lr": 1e-3}, {"params": ours_p[3:], "lr": 1e-4, "weight_decay": 0.1}],
                         max_grad_norm=1.0)
    ref = torch.optim.AdamW([{"params": ref_p[:3], "lr": 1e-3}, {"params": ref_p[3:], "lr": 1e-4, "weight_decay": 0.1}])
    g = torch.Generator().manual_seed(4)

    def run(o, r, op, rp, n):
        for _ in range(n):
            for a, b in zip(op, rp):
                gr = torch.randn(a.shape, generator=g).to(dev)
                a.grad, b.grad = gr.clone(), gr.clone()
            torch.nn.utils.clip_grad_norm_(rp, max_norm=1.0)
            r.step()
            o.step()

    run(ours, ref, ours_p, ref_p, 3)
    # resume: the reference optimizer's state_dict goes into a fresh FusedAdamW (checkpoint interchange)
    ours2_p = [torch.nn.Parameter(p.detach().clone()) for p in ref_p]
    ours2 = tt.FusedAdamW([{"params": ours2_p[:3]}, {"params": ours2_p[3:]}], max_grad_norm=1.0)
    import copy
    ours2.load_state_dict(copy.deepcopy(ref.state_dict()))      # (a checkpoint file would not share tensors either)
    run(ours2, ref, ours2_p, ref_p, 3)
    for a, b in zip(ours2_p, ref_p):
        assert torch.allclose(a, b, rtol=3e-6, atol=3e-7), (a.shape, (a - b).abs().max().item())


@pytest.mark.gpu
def test_fused_adam_trains_a_model_like_torch_adam():
    """Whole path: model fwd/bwd through the library, then clip + AdamW fused vs the reference's two calls."""
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m1 = tt.SimpleTwoTower(12, 8).to(dev).train()
    m2 = tt.SimpleTwoTower(12, 8).to(dev).train()
    m2.load_state_dict(m1.state_dict())
    for m in (m1, m2):
        for mod in m.modules():
            if isinstance(mod, tt.GRU):
                mod.dropout = 0.0
            if isinstance(mod, torch.nn.Dropout):
                mod.p = 0.0
    o1 = tt.FusedAdamW(m1.parameters(), lr=1e-3, weight_decay=0.01, max_grad_norm=1.0)
    o2 = torch.optim.AdamW(m2.parameters(), lr=1e-3, weight_decay=0.01)
    crit = tt.SymmetricInfoNCELoss(0.1)
    g = torch.Generator().manual_seed(1)
    for _ in range(4):
        q, d = torch.randn(16, 5, 12, generator=g).to(dev), torch.randn(16, 7, 12, generator=g).to(dev)
        for m, o in ((m1, o1), (m2, o2)):
            o.zero_grad()
            loss = crit(*m(q, d))
            loss.backward()
            if o is o2:
                torch.nn.utils.clip_grad_norm_(m.parameters(), max_norm=1.0)
            o.step()
    for (k, a), (_, b) in zip(m1.state_dict().items(), m2.state_dict().items()):
        assert rel_err(a, b) < 1e-5, k


@pytest.mark.gpu
def test_pinned_id_ring_delivers_tokenised_batches(vocab):
    dev = torch.device("cuda:0")
    rng = random.Random(3)
    queries = [" ".join(rng.choice(G["words"] + ["oov"]) for _ in range(rng.randint(0, 12))) for _ in range(70)]
    docs = [" ".join(rng.choice(G["words"] + ["oov"]) for _ in range(rng.randint(0, 40))) for _ in range(70)]
    seen_q, seen_d = [], []
    for q, d in tt.id_batches(vocab, queries, docs, batch_size=16, query_len=8, doc_len=24, device=dev):
        assert q.is_cuda and q.dtype == torch.int64 and q.shape[1] == 8 and d.shape[1] == 24
        seen_q.append(q.cpu().clone())
        seen_d.append(d.cpu().clone())
    assert torch.cat(seen_q).tolist() == [rpl.simple_ids(t, INDEX, 8) for t in queries]
    assert torch.cat(seen_d).tolist() == [rpl.simple_ids(t, INDEX, 24) for t in docs]
    assert [len(x) for x in seen_q] == [16, 16, 16, 16, 6]
    assert len(list(tt.id_batches(vocab, queries, docs, 16, 8, 24, dev, drop_last=True))) == 4


@pytest.mark.gpu
def test_search_service_matches_oracle_and_cache_roundtrip(vocab, tmp_path):
    from oracle import ref_port as rp
    dev = torch.device("cuda:0")
    E, H, L = 10, 8, 12
    torch.manual_seed(0)
    port = rp.build_port("margin", E, H).eval()
    model = tt.TwoTowerModel(E, H)
    model.load_state_dict(port.state_dict())
    table = torch.randn(vocab.table_rows(), E) * 0.5
    table[0] = 0
    model.to(dev).eval().attach_embeddings(table.to(dev), pad_id=0)
    rng = random.Random(9)
    docs = [" ".join(rng.choice(G["words"]) for _ in range(rng.randint(1, 20))) + " tail" * (i % 3 * 40) for i in range(50)]
    cache = str(tmp_path / "doc_embeddings.pt")
    truth = {"what is ai": [docs[7]]}
    svc = tt.SearchService(model, vocab, docs, max_length=L, top_k=3, cache_path=cache, query_to_docs=truth)
    assert os.path.exists(cache)
    cached = torch.load(cache, map_location="cpu")                        # bare tensor, as app.py:65 writes it
    assert torch.is_tensor(cached) and cached.shape == (50, H)
    # oracle: the reference flow of app.py:55-62,88-101 on CPU with the port
    with torch.no_grad():
        doc_ids = torch.tensor([rpl.simple_ids(t, INDEX, L) for t in docs])
        ref_docs = torch.cat([port.encode_doc(table[doc_ids[i:i + 1]]) for i in range(len(docs))], 0)
        assert rel_err(cached, ref_docs) < 1e-3
        for query in ["what is ai", "how to cook pasta", "zzz", ""]:
            qv = port.encode_query(table[torch.tensor([rpl.simple_ids(query, INDEX, L)])])
            sims = torch.nn.functional.cosine_similarity(qv.unsqueeze(1), ref_docs.unsqueeze(0), dim=2).squeeze()
            top = rp.stable_topk(sims.unsqueeze(0), 3)[1]
            out = svc.search(query)
            assert out["query"] == query and [r["rank"] for r in out["results"]] == [1, 2, 3]
            cut = lambda x: x[:200] + "..." if len(x) > 200 else x
            # indices must agree unless two of the leading scores are closer than the fp32 tolerance of the encoders
            gap = sims.sort(descending=True).values[:4].diff().abs().min().item()
            if gap > 5e-3:
                for r, i in zip(out["results"], top[0].tolist()):
                    assert r["text"] == cut(docs[i])
                    assert abs(r["score"] - sims[i].item()) < 2e-3
                    assert r["is_ground_truth"] == (docs[i] in truth.get(query, []))
            else:
                for r in out["results"]:
                    assert abs(r["score"] - sims.max().item()) < 0.5
    # a second service instance starts from the cache file and returns the same answers
    svc2 = tt.SearchService(model, vocab, docs, max_length=L, top_k=3, cache_path=cache)
    assert torch.equal(svc2.doc_embeddings.cpu(), cached)
    assert svc2.search("what is ai")["results"][0]["text"] == svc.search("what is ai")["results"][0]["text"]


@pytest.mark.gpu
def test_compute_mrr_matches_oracle_flow(vocab):
    """validate_margin.compute_mrr through the package: margin text preprocessing -> ids -> device towers -> full ranking.
    Oracle: the port on CPU with oracle/ref_pipeline.margin_ids + full_rank_rr (pinned to the live reference on CPU in
    tests/test_full_rank_mrr.py).  Only (query, relevant document) pairs whose relevant score is > 1e-2 away from every other
    candidate's are compared (11 of the 15 with these seeds): the rank cannot flip inside the encoders' 1e-3 tolerance."""
    from oracle import ref_port as rp
    dev = torch.device("cuda:0")
    E, H, L = 10, 8, 30
    torch.manual_seed(1)
    port = rp.build_port("margin", E, H).eval()
    model = tt.TwoTowerModel(E, H)
    model.load_state_dict(port.state_dict())
    table = torch.randn(max(INDEX.values()) + 1, E, generator=torch.Generator().manual_seed(7)) * 0.5
    table[0] = 0
    assert vocab.table_rows() == table.shape[0]
    model.to(dev).train().attach_embeddings(table.to(dev), pad_id=0)
    rng = random.Random(4)
    cands = [" ".join(rng.choice(G["words"]) for _ in range(rng.randint(1, 25))) for _ in range(12)]
    queries = [cands[3], "what is a " + cands[7], " ".join(rng.choice(G["words"]) for _ in range(6)), "zzz", cands[10] + " the"]
    checked = 0
    with torch.no_grad():
        dv = port.encode_doc(table[torch.tensor([rpl.margin_ids(c, INDEX, L) for c in cands])])
    for query in queries:
        with torch.no_grad():
            qv = port.encode_query(table[torch.tensor([rpl.margin_ids(query, INDEX, L)])])
            sims = torch.nn.functional.cosine_similarity(qv.unsqueeze(1), dv.unsqueeze(0), dim=2).squeeze(0)
        for j in (0, 6, 9):
            relevant = {cands[j]}
            rr, top = tt.compute_mrr(model, query, relevant, cands, vocab, max_length=L, batch_size=5)
            assert model.training                                          # mode restored
            assert len(top) == 3 and all(isinstance(d, str) for d, _ in top)
            want_rr, want_top = rpl.full_rank_rr(sims.tolist(), cands, relevant)
            others = torch.cat([sims[:j], sims[j + 1:]])
            if (others - sims[j]).abs().min().item() > 1e-2:
                assert rr == want_rr
                checked += 1
            assert abs(top[0][1] - want_top[0][1]) < 5e-3
    assert checked >= 8
    assert tt.compute_mrr(model, "anything", {"x"}, [], vocab) == (0.0, [])


@pytest.mark.gpu
def test_fused_adam_skips_a_step_with_non_finite_gradients():
    """An overflow of the fp16 gradient planes (inf / NaN weight gradients) must not reach the weights: the step is skipped
    and counted, as torch.cuda.amp.GradScaler does; the next finite step continues from the untouched state."""
    dev = torch.device("cuda:0")
    ours_p, ref_p = _rand_params(dev, 5, SHAPES), _rand_params(dev, 5, SHAPES)
    for capturable in (False, True):
        ours = tt.FusedAdam(ours_p, lr=1e-3, max_grad_norm=1.0, capturable=capturable)
        ref = torch.optim.Adam(ref_p, lr=1e-3)
        g = torch.Generator().manual_seed(6)
        for step in range(4):
            bad = step == 1
            for a, b in zip(ours_p, ref_p):
                gr = torch.randn(a.shape, generator=g).to(dev)
                a.grad, b.grad = gr.clone(), gr.clone()
            if bad:
                ours_p[2].grad[5] = float("inf")
                before = [p.detach().clone() for p in ours_p]
            else:
                torch.nn.utils.clip_grad_norm_(ref_p, max_norm=1.0)
                ref.step()
            ours.step()
            if bad:
                assert all(torch.equal(a, b) for a, b in zip(ours_p, before))
                assert float(ours.skipped_steps) == 1.0
        if capturable:          # the device step count did not advance on the skipped step: bit-for-bit torch's trajectory
            assert float(ours.state[ours_p[0]]["step"]) == 3.0
            for a, b in zip(ours_p, ref_p):
                assert torch.allclose(a, b, rtol=2e-6, atol=2e-7)
        else:                   # host step count advanced (documented): only the bias correction differs, slightly
            for a, b in zip(ours_p, ref_p):
                assert torch.allclose(a, b, rtol=0.3, atol=2e-3)
        with torch.no_grad():
            for a, b in zip(ours_p, ref_p):
                b.copy_(a)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["enhanced", "simple"])
def test_graphed_train_step_equals_the_eager_loop(kind):
    """N4: the whole step (zero_grad, both towers, loss, backward, clip + AdamW, hard-negative mining) captured in one CUDA
    graph and replayed equals the eager loop step for step (dropout off: the masks come from different seeds)."""
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    mk = (lambda: tt.EnhancedTwoTowerModel(20, 32)) if kind == "enhanced" else (lambda: tt.SimpleTwoTower(20, 64))
    m1, m2 = mk().to(dev).train(), mk().to(dev).train()
    m2.load_state_dict(m1.state_dict())
    table = (torch.randn(500, 20) * 0.3).to(dev)
    for m in (m1, m2):
        m.attach_embeddings(table, pad_id=0)
        for mod in m.modules():
            if isinstance(mod, tt.GRU):
                mod.dropout = 0.0
            if isinstance(mod, torch.nn.Dropout):
                mod.p = 0.0
    crit = tt.MarginRankingLoss() if kind == "enhanced" else tt.SymmetricInfoNCELoss(0.1)
    o1 = tt.FusedAdamW(m1.parameters(), lr=1e-3, weight_decay=0.01, max_grad_norm=1.0, capturable=True)
    o2 = tt.FusedAdamW(m2.parameters(), lr=1e-3, weight_decay=0.01, max_grad_norm=1.0)
    step = tt.GraphedTrainStep(m1, crit, o1, mine_k=4)
    g = torch.Generator().manual_seed(1)
    for it in range(5):
        B = 48 if it != 3 else 40                     # a second shape in the middle: second capture, shared state
        q = torch.randint(0, 500, (B, 6), generator=g).to(dev)
        d = torch.randint(0, 500, (B, 11), generator=g).to(dev)
        loss1, hard1 = step(q, d)
        loss1, hard1 = loss1.clone(), hard1.clone()
        o2.zero_grad(set_to_none=True)
        qv, dv = m2(q, d)
        loss2 = crit(qv, dv)
        loss2.backward()
        o2.step()
        hard2 = tt.get_hard_negatives_batch(qv.detach(), dv.detach(), k=4)
        assert abs(loss1.item() - loss2.item()) <= 1e-5 * abs(loss2.item()), (it, loss1.item(), loss2.item())
        assert torch.equal(hard1, hard2), it
        for (k, a), (_, b) in zip(m1.state_dict().items(), m2.state_dict().items()):
            assert rel_err(a, b) < 1e-5, (it, k)
    assert step.replays == 5 and len(step._graphs) == 2
    assert step.launches_per_replay(q, d) > 20


@pytest.mark.gpu
def test_graphed_train_step_draws_new_dropout_masks_on_every_replay():
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m = tt.SimpleTwoTower(20, 64).to(dev).train()       # GRU inter-layer dropout 0.1 + head dropout 0.1
    o = tt.FusedAdamW(m.parameters(), lr=0.0, weight_decay=0.0, capturable=True)     # lr 0: weights stay put
    step = tt.GraphedTrainStep(m, tt.SymmetricInfoNCELoss(0.1), o)
    q, d = torch.randn(32, 5, 20, device=dev), torch.randn(32, 7, 20, device=dev)
    losses = [step(q, d)[0].item() for _ in range(4)]
    assert len(set(losses)) == 4, losses                # same inputs, same weights: only the masks differ


@pytest.mark.gpu
def test_graphed_towers_equal_the_eager_step_and_fire_grad_hooks():
    """The data-parallel form of the graphed step: tower forward / backward replayed from two CUDA graphs, loss / optimizer
    eager.  Equal to the eager loop step for step (two shapes), and post-accumulate-grad hooks (the all-reducer's) fire."""
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m1, m2 = tt.EnhancedTwoTowerModel(20, 32).to(dev).train(), tt.EnhancedTwoTowerModel(20, 32).to(dev).train()
    m2.load_state_dict(m1.state_dict())
    table = (torch.randn(500, 20) * 0.3).to(dev)
    for m in (m1, m2):
        m.attach_embeddings(table, pad_id=0)
        for mod in m.modules():
            if isinstance(mod, tt.GRU):
                mod.dropout = 0.0
            if isinstance(mod, torch.nn.Dropout):
                mod.p = 0.0
    crit = tt.MarginRankingLoss()
    o1 = tt.FusedAdamW(m1.parameters(), lr=1e-3, weight_decay=0.01, max_grad_norm=1.0)
    o2 = tt.FusedAdamW(m2.parameters(), lr=1e-3, weight_decay=0.01, max_grad_norm=1.0)
    towers = tt.GraphedTowers(m1)
    fired = []
    hooks = [p.register_post_accumulate_grad_hook(lambda p: fired.append(1)) for p in m1.parameters()]
    g = torch.Generator().manual_seed(1)
    for it in range(5):
        B = 48 if it != 3 else 40
        q = torch.randint(0, 500, (B, 6), generator=g).to(dev)
        d = torch.randint(0, 500, (B, 11), generator=g).to(dev)
        fired.clear()
        o1.zero_grad(set_to_none=True)
        qv1, dv1 = towers(q, d)
        loss1 = crit(qv1, dv1)
        loss1.backward()
        o1.step()
        assert len(fired) == len(list(m1.parameters())), it
        o2.zero_grad(set_to_none=True)
        qv2, dv2 = m2(q, d)
        loss2 = crit(qv2, dv2)
        loss2.backward()
        o2.step()
        assert abs(loss1.item() - loss2.item()) <= 1e-5 * abs(loss2.item()), (it, loss1.item(), loss2.item())
        for (k, a), (_, b) in zip(m1.state_dict().items(), m2.state_dict().items()):
            assert rel_err(a, b) < 1e-5, (it, k)
    for h in hooks:
        h.remove()
    assert len(towers._runs) == 2 and towers.launches_per_replay(q, d) > 20


@pytest.mark.gpu
def test_graphed_towers_draw_new_dropout_masks_on_every_call():
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    m = tt.SimpleTwoTower(20, 64).to(dev).train()
    towers = tt.GraphedTowers(m)
    q, d = torch.randn(32, 5, 20, device=dev), torch.randn(32, 7, 20, device=dev)
    crit = tt.SymmetricInfoNCELoss(0.1)
    losses = []
    for _ in range(4):
        qv, dv = towers(q, d)
        loss = crit(qv, dv)
        loss.backward()                                  # backward of the same call sees the same masks: finite gradients
        assert all(torch.isfinite(p.grad).all() for p in m.parameters())
        m.zero_grad(set_to_none=True)
        losses.append(loss.item())
    assert len(set(losses)) == 4, losses
