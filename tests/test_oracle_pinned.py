"""CPU: pin the oracle (oracle/ref_port.py, oracle/ref_numpy.py) against the golden vectors produced by the
UNMODIFIED reference (tests/golden/*.npz, made by oracle/gen_golden.py) and, when /root/reference is present,
against the reference itself."""
import numpy as np
import pytest
import torch

from conftest import GOLDEN_MODELS, golden_state_dict, load_golden, rel_err
from oracle import ref_numpy as rn
from oracle import ref_port as rp
from oracle.ref_import import load_reference, reference_available

LOSSES = {
    "infonce_sym_t0.1": lambda q, d: rp.info_nce_symmetric(q, d, 0.1),
    "infonce_q2d_t0.1": lambda q, d: rp.info_nce_q2d(q, d, 0.1),
    "infonce_q2d_t0.07": lambda q, d: rp.info_nce_q2d(q, d, 0.07),
    "margin_ce_m0.2_t0.1": lambda q, d: rp.margin_ce_inbatch(q, d, 0.2, 0.1),
}


def _port(name):
    g = load_golden(name)
    m = rp.build_port(GOLDEN_MODELS[name], int(g["E"]), int(g["H"]))
    m.load_state_dict(golden_state_dict(g), strict=True)
    return g, m


@pytest.mark.parametrize("name", list(GOLDEN_MODELS))
def test_port_forward_matches_golden(name):
    g, m = _port(name)
    m.eval()
    q, d = torch.from_numpy(g["q"]), torch.from_numpy(g["d"])
    with torch.no_grad():
        assert rel_err(m.encode_query(q), g["eval/encode_query"]) < 1e-6
        assert rel_err(m.encode_doc(d), g["eval/encode_doc"]) < 1e-6
        out = m(q, d)
    if isinstance(out, tuple):
        assert rel_err(out[0], g["eval/forward_q"]) < 1e-6 and rel_err(out[1], g["eval/forward_d"]) < 1e-6
    else:
        assert rel_err(out, g["eval/forward_sim"]) < 1e-6


@pytest.mark.parametrize("name", list(GOLDEN_MODELS))
def test_port_losses_and_grads_match_golden(name):
    g, m = _port(name)
    rp.disable_dropout(m).train()
    q, d = torch.from_numpy(g["q"]), torch.from_numpy(g["d"])
    tags = sorted({k.split("/")[1] for k in g if k.startswith("train/")})
    assert tags
    for tag in tags:
        m.zero_grad()
        qv, dv = m(q, d)
        loss = LOSSES[tag](qv, dv)
        loss.backward()
        assert abs(loss.item() - float(g[f"train/{tag}/loss"])) <= 1e-5 * max(1.0, abs(loss.item()))
        for k, p in m.named_parameters():
            ref = torch.from_numpy(g[f"train/{tag}/grad/{k}"])
            got = p.grad if p.grad is not None else torch.zeros_like(p)
            assert (got - ref).norm() <= 1e-5 * ref.norm() + 1e-7, (tag, k)


@pytest.mark.parametrize("name", list(GOLDEN_MODELS))
def test_numpy_restatement_matches_golden(name):
    g = load_golden(name)
    kind = GOLDEN_MODELS[name]
    params = {k[3:]: v for k, v in g.items() if k.startswith("sd/")}
    hq, hd = ("projection", "projection") if kind == "margin" else ("query_proj", "doc_proj")
    qv = rn.tower(g["q"], params, "query_encoder", hq, kind)
    dv = rn.tower(g["d"], params, "doc_encoder", hd, kind)
    assert rel_err(qv, g["eval/encode_query"]) < 2e-6
    assert rel_err(dv, g["eval/encode_doc"]) < 2e-6


def test_numpy_losses_match_golden():
    g = load_golden("losses_mining")
    q, p, neg = (g[k].astype(np.float64) for k in ("q", "pos", "neg"))
    B = q.shape[0]
    qn, pn = rn.l2_normalize(q), rn.l2_normalize(p)
    sym = 0.5 * (rn.score_ce(qn, pn, 10.0) + rn.score_ce(pn, qn, 10.0))
    assert abs(sym - g["simple_infonce_t0.1/loss"]) < 1e-5
    assert abs(rn.score_ce(q, p, 1 / 0.07, normalize=True) - g["margin_infonce_t0.07/loss"]) < 1e-5
    assert abs(rn.score_ce(q, p, 1 / 0.07, normalize=True) - g["enhanced_infonce_t0.07/loss"]) < 1e-5
    assert abs(rn.score_ce(q, p, 10.0, margin=0.2) - g["margin_ce_m0.2_t0.1/loss"]) < 1e-4
    assert abs(rn.score_ce(q, p, 20.0, margin=0.5) - g["margin_ce_m0.5_t0.05/loss"]) < 1e-4
    assert abs(rn.hinge_explicit_negatives(q, p, neg, 0.2) - g["hinge_m0.2/loss"]) < 1e-6
    assert abs(rn.hinge_explicit_negatives(q, p, neg, 1.0) - g["hinge_m1.0/loss"]) < 1e-6
    assert B == 9


def test_port_losses_match_golden_with_grads():
    g = load_golden("losses_mining")
    for tag, fn, names in (
        ("margin_infonce_t0.07", lambda q, p, n: rp.info_nce_q2d(q, p, 0.07), ("dq", "dpos")),
        ("margin_ce_m0.2_t0.1", lambda q, p, n: rp.margin_ce_inbatch(q, p, 0.2, 0.1), ("dq", "dpos")),
        ("hinge_m0.2", lambda q, p, n: rp.hinge_explicit_negatives(q, p, n, 0.2), ("dq", "dpos", "dneg")),
    ):
        q, p, n = (torch.from_numpy(g[k]).clone().requires_grad_(True) for k in ("q", "pos", "neg"))
        loss = fn(q, p, n)
        grads = torch.autograd.grad(loss, (q, p, n)[: len(names)])
        assert abs(loss.item() - float(g[f"{tag}/loss"])) < 1e-5
        for nm, gr in zip(names, grads):
            assert rel_err(gr, g[f"{tag}/{nm}"]) < 1e-5


def test_mining_and_search_match_golden():
    g = load_golden("losses_mining")
    q, docs = torch.from_numpy(g["q"]), torch.from_numpy(g["docs"])
    off = int(g["hardneg_pos_offset"])
    for k in (5, 8):
        assert np.array_equal(rp.hard_negatives_rows(q, docs, off, k).numpy(), g[f"hardneg_k{k}"])
        assert np.array_equal(rn.hard_negatives_rows(g["q"].astype(np.float64), g["docs"].astype(np.float64), off, k),
                              g[f"hardneg_k{k}"])
    val, idx = rp.search_topk(q, docs, 10)
    assert np.array_equal(idx.numpy(), g["search_top10_idx"])
    assert np.allclose(val.numpy(), g["search_top10_val"], atol=1e-6)


def test_stable_topk_breaks_ties_to_lowest_index():
    s = torch.tensor([0.5, 0.9, 0.9, 0.1, 0.9, 0.5])
    assert rp.stable_topk(s, 4)[1].tolist() == [1, 2, 4, 0]
    assert rn.stable_topk(s.numpy(), 4)[1].tolist() == [1, 2, 4, 0]


def test_mrr_recall_definition():
    top = torch.tensor([[5, 2, 9], [1, 3, 4], [7, 8, 6]])
    out = rp.mrr_recall(top, torch.tensor([2, 1, 0]), ks=(1, 3))
    assert abs(out["MRR"] - (0.5 + 1.0 + 0.0) / 3) < 1e-12
    assert abs(out["R@1"] - 1 / 3) < 1e-12 and abs(out["R@3"] - 2 / 3) < 1e-12


def test_lengths_semantics_match_pack_padded_sequence():
    torch.manual_seed(3)
    gru = torch.nn.GRU(5, 4, num_layers=2, batch_first=True, bidirectional=True)
    x = torch.randn(3, 6, 5)
    lens = [6, 2, 4]
    packed = torch.nn.utils.rnn.pack_padded_sequence(x, lens, batch_first=True, enforce_sorted=False)
    out_p, h_n = gru(packed)
    out_ref, _ = torch.nn.utils.rnn.pad_packed_sequence(out_p, batch_first=True, total_length=6)
    params = {"g." + k: v.detach().numpy() for k, v in gru.named_parameters()}
    out, hn = rn.gru_stack(x.numpy(), params, "g", lengths=lens)
    assert rel_err(out, out_ref.detach()) < 1e-6 and rel_err(hn, h_n.detach()) < 1e-6


@pytest.mark.skipif(not reference_available(), reason="/root/reference not present (GPU box)")
@pytest.mark.parametrize("kind,E,H", [("simple", 300, 16), ("margin", 300, 16), ("enhanced", 300, 8)])
def test_port_equals_live_reference(kind, E, H):
    S, M, En = load_reference()
    cls = {"simple": S.SimpleTwoTower, "margin": M.TwoTowerModel, "enhanced": En.EnhancedTwoTowerModel}[kind]
    torch.manual_seed(5)
    ref = cls(E, H)
    torch.manual_seed(5)
    port = rp.build_port(kind, E, H)          # same construction order -> same init from the same seed
    for (k1, v1), (k2, v2) in zip(ref.state_dict().items(), port.state_dict().items()):
        assert k1 == k2 and torch.equal(v1, v2)
    ref.eval(); port.eval()
    q, d = torch.randn(4, 5, E), torch.randn(4, 9, E)
    with torch.no_grad():
        a, b = ref(q, d), port(q, d)
    if isinstance(a, tuple):
        assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    else:
        assert torch.equal(a, b)
