"""Row a16: full-rank MRR of validate_margin.compute_mrr.  CPU: the ranking semantics (cosine, stable descending
sort, first relevant hit, three best) against the oracle restatement and - where /root/reference exists - against the
unmodified reference function run on its own model."""
import random

import numpy as np
import pytest
import torch

from oracle import ref_pipeline as rpl
from oracle.ref_import import load_reference, reference_available

import two_towers_b200 as tt


def test_first_relevant_rank_equals_oracle_with_ties_and_misses():
    rng = np.random.default_rng(0)
    Q, N, D = 40, 57, 12
    docs = rng.standard_normal((N, D)).astype(np.float32)
    docs[20] = docs[3]; docs[41] = docs[3]; docs[8] = 2.0 * docs[30]          # exact ties (cosine ignores the scale)
    q = rng.standard_normal((Q, D)).astype(np.float32)
    q[0] = docs[3]
    rel = rng.random((Q, N)) < 0.05
    rel[1] = False                                                            # a query without relevant candidates
    rel[0] = False; rel[0, 41] = True                                         # relevant doc is the LAST of three tied ones
    rr, idx, val = tt.first_relevant_rank(torch.from_numpy(q), torch.from_numpy(docs), torch.from_numpy(rel))
    sims = torch.nn.functional.cosine_similarity(torch.from_numpy(q).unsqueeze(1), torch.from_numpy(docs).unsqueeze(0), dim=2)
    names = [f"doc{j}" for j in range(N)]
    for i in range(Q):
        want_rr, want_top = rpl.full_rank_rr(sims[i].tolist(), names, {names[j] for j in np.nonzero(rel[i])[0]})
        assert rr[i].item() == want_rr
        assert [names[j] for j in idx[i].tolist()] == [d for d, _ in want_top]
        assert val[i].tolist() == [s for _, s in want_top]
    assert rr[1].item() == 0.0 and rr[0].item() == 1.0 / 3.0
    assert idx[0].tolist() == [3, 20, 41]


def test_first_relevant_rank_shapes():
    q = torch.randn(5, 6)
    rr, idx, val = tt.first_relevant_rank(q[0], torch.randn(2, 6), torch.tensor([[False, True]]))
    assert rr.shape == (1,) and idx.shape == (1, 2) and rr.item() in (1.0, 0.5)
    per_query = torch.randn(5, 9, 6)                                          # one candidate list per query
    rel = torch.zeros(5, 9, dtype=torch.bool)
    rel[:, 4] = True
    rr, idx, val = tt.first_relevant_rank(q, per_query, rel)
    sims = torch.nn.functional.cosine_similarity(q.unsqueeze(1), per_query, dim=2)
    want = 1.0 / (1 + (sims > sims[:, 4:5]).sum(1)).double()
    assert torch.equal(rr, want)


@pytest.mark.skipif(not reference_available(), reason="reference tree not present")
def test_first_relevant_rank_equals_live_reference_compute_mrr():
    import importlib
    _, margin, _ = load_reference()
    vm = importlib.import_module("validate_margin")

    class W2V(dict):
        vector_size = 300

    rng = random.Random(5)
    words = [f"w{i}" for i in range(60)] + ["is", "a", "has", "part", "of", "controls"]
    npr = np.random.default_rng(1)
    w2v = W2V({w: npr.standard_normal(300).astype(np.float32) * 0.3 for w in words})
    torch.manual_seed(3)
    model = margin.TwoTowerModel(300, 16).eval()
    cands = [" ".join(rng.choice(words + ["oov"]) for _ in range(rng.randint(1, 25))) for _ in range(40)]
    cands[17] = cands[5]; cands[30] = cands[5]                                # duplicate texts: exact score ties
    compared = 0
    for query, relevant in (("w1 w2 is a w3", {cands[9], cands[22]}), (cands[5], {cands[30]}), ("w7 oov", set()),
                            ("w4 w4 w4", {cands[0]})):
        ref_rr, ref_top = vm.compute_mrr(model, query, relevant, cands, w2v, torch.device("cpu"))
        with torch.no_grad():
            qv = model.encode_query(margin.SimpleDataset.text_to_embedding(query, w2v).unsqueeze(0))
            dv = torch.cat([model.encode_doc(margin.SimpleDataset.text_to_embedding(c, w2v).unsqueeze(0)) for c in cands], 0)
        mask = torch.tensor([[c in relevant for c in cands]])
        rr, idx, val = tt.first_relevant_rank(qv, dv, mask)
        # (the reference encodes candidates 128 at a time, here one at a time: scores agree to fp32 rounding)
        gaps_ok = torch.sort(torch.nn.functional.cosine_similarity(qv, dv), descending=True).values.diff().abs()
        if (gaps_ok[gaps_ok > 0] > 1e-5).all():
            assert rr.item() == ref_rr
            assert [cands[i] for i in idx[0].tolist()] == [d for d, _ in ref_top]
            assert np.allclose(val[0].tolist(), [s for _, s in ref_top], atol=1e-5)
            compared += 1
    assert compared >= 3


def test_compute_mrr_glue_on_cpu_with_the_port_as_model():
    """compute_mrr's own logic (margin tokenisation, batching, relevance by text, mode restore) with the oracle port standing
    in for the device model: must reproduce oracle margin_ids + full_rank_rr exactly (same arithmetic on both sides)."""
    import json
    import os
    from conftest import GOLDEN
    from oracle import ref_port as rp
    G = json.load(open(os.path.join(GOLDEN, "pipeline.json")))
    vocab = tt.Vocabulary(G["words"], pad_id=0, first_id=1)
    index = {}
    for i, w in enumerate(G["words"]):
        index.setdefault(w, i + 1)                                            # duplicate words keep their first row
    E, H, L = 10, 8, 30
    torch.manual_seed(1)
    port = rp.build_port("margin", E, H).train()
    table = torch.randn(vocab.table_rows(), E, generator=torch.Generator().manual_seed(7)) * 0.5
    table[0] = 0

    class IdModel(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.port = port

        def encode_query(self, ids):
            return self.port.encode_query(table[ids])

        def encode_doc(self, ids):
            return self.port.encode_doc(table[ids])

    model = IdModel().train()
    rng = random.Random(4)
    cands = [" ".join(rng.choice(G["words"]) for _ in range(rng.randint(1, 25))) for _ in range(12)]
    cands[8] = cands[1]                                                       # duplicate text: tie, both relevant by text
    for query, relevant in ((cands[3], {cands[9]}), ("what is a " + cands[7], {cands[1]}), ("zzz", {"not a candidate"})):
        rr, top = tt.compute_mrr(model, query, relevant, cands, vocab, max_length=L, batch_size=5)
        assert model.training
        port.eval()
        with torch.no_grad():
            qv = port.encode_query(table[torch.tensor([rpl.margin_ids(query, index, L)])])
            dv = torch.cat([port.encode_doc(table[torch.tensor([rpl.margin_ids(c, index, L) for c in cands[i:i + 5]])])
                            for i in range(0, 12, 5)], 0)
            sims = torch.nn.functional.cosine_similarity(qv.unsqueeze(1), dv.unsqueeze(0), dim=2).squeeze(0)
        port.train()
        want_rr, want_top = rpl.full_rank_rr(sims.tolist(), cands, relevant)
        assert rr == want_rr and [d for d, _ in top] == [d for d, _ in want_top]
        assert [s for _, s in top] == [s for _, s in want_top]
    assert tt.compute_mrr(model, "anything", {"x"}, [], vocab) == (0.0, [])
