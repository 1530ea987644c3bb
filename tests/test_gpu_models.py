"""GPU: drop-in parity of the three models + losses against (a) the golden vectors made by the unmodified
reference and (b) the CPU port at the BASELINE configs' layer sizes.  Bound (north_star): embeddings, loss and
gradients <= 1e-3 relative, fp32 accumulation; exact-fp32 engine is expected near 1e-5."""
import pytest
import torch

import two_towers_b200 as tt
from conftest import GOLDEN_MODELS, golden_state_dict, load_golden, rel_err
from oracle import ref_port as rp
from two_towers_b200 import simple_two_tower as S

pytestmark = pytest.mark.gpu
DEV = "cuda"
TOL = 1e-3
CLS = {"simple": tt.SimpleTwoTower, "margin": tt.TwoTowerModel, "enhanced": tt.EnhancedTwoTowerModel}
LOSSES = {
    "infonce_sym_t0.1": lambda: S.InfoNCELoss(0.1),
    "infonce_q2d_t0.1": lambda: tt.InfoNCELoss(0.1),
    "infonce_q2d_t0.07": lambda: tt.InfoNCELoss(),
    "margin_ce_m0.2_t0.1": lambda: tt.MarginRankingLoss(),
}
PORT_LOSSES = {
    "infonce_sym_t0.1": lambda q, d: rp.info_nce_symmetric(q, d, 0.1),
    "infonce_q2d_t0.1": lambda q, d: rp.info_nce_q2d(q, d, 0.1),
    "infonce_q2d_t0.07": lambda q, d: rp.info_nce_q2d(q, d, 0.07),
    "margin_ce_m0.2_t0.1": lambda q, d: rp.margin_ce_inbatch(q, d, 0.2, 0.1),
}


def _no_dropout(m):
    for mod in m.modules():
        if isinstance(mod, torch.nn.Dropout):
            mod.p = 0.0
        if isinstance(mod, tt.GRU):
            mod.dropout = 0.0
    return m


def _grad_close(got, ref, scale):
    """relative error, with an absolute floor for analytically-zero gradients (e.g. the last bias under CE)."""
    got, ref = got.double().cpu(), torch.as_tensor(ref).double()
    return (got - ref).norm().item() <= TOL * ref.norm().item() + 1e-6 * scale


@pytest.mark.parametrize("name", list(GOLDEN_MODELS))
def test_model_matches_reference_golden(name):
    g = load_golden(name)
    m = CLS[GOLDEN_MODELS[name]](int(g["E"]), int(g["H"]))
    m.load_state_dict(golden_state_dict(g), strict=True)
    m.to(DEV).eval()
    q, d = torch.from_numpy(g["q"]).to(DEV), torch.from_numpy(g["d"]).to(DEV)
    with torch.no_grad():
        assert rel_err(m.encode_query(q), g["eval/encode_query"]) < TOL
        assert rel_err(m.encode_doc(d), g["eval/encode_doc"]) < TOL
        out = m(q, d)
    if isinstance(out, tuple):
        assert rel_err(out[0], g["eval/forward_q"]) < TOL and rel_err(out[1], g["eval/forward_d"]) < TOL
    else:
        assert out.shape == (q.shape[0], q.shape[0]) and rel_err(out, g["eval/forward_sim"]) < TOL
    _no_dropout(m).train()
    tags = sorted({k.split("/")[1] for k in g if k.startswith("train/")})
    for tag in tags:
        m.zero_grad()
        qv, dv = m(q, d)
        assert rel_err(qv, g[f"train/{tag}/q"]) < TOL and rel_err(dv, g[f"train/{tag}/d"]) < TOL
        loss = LOSSES[tag]()(qv, dv)
        ref_loss = float(g[f"train/{tag}/loss"])
        assert abs(loss.item() - ref_loss) <= TOL * abs(ref_loss), tag
        loss.backward()
        scale = max(float(torch.as_tensor(g[f"train/{tag}/grad/{k}"]).norm()) for k, _ in m.named_parameters())
        for k, p in m.named_parameters():
            assert p.grad is not None, k
            assert _grad_close(p.grad, g[f"train/{tag}/grad/{k}"], scale), (tag, k, rel_err(p.grad, g[f"train/{tag}/grad/{k}"]))


CONFIG_CASES = [
    # kind, E, hidden_dim, B, Tq, Td, loss tag   (layer sizes of BASELINE configs C1', C2, C3; batch cut to CPU-oracle size)
    ("simple", 300, 128, 48, 32, 32, "infonce_sym_t0.1"),
    ("margin", 300, 256, 24, 32, 128, "margin_ce_m0.2_t0.1"),
    ("margin", 300, 256, 24, 32, 128, "infonce_q2d_t0.1"),
    ("enhanced", 300, 128, 24, 32, 128, "margin_ce_m0.2_t0.1"),
    ("enhanced", 300, 128, 24, 32, 128, "infonce_q2d_t0.07"),
]


@pytest.mark.parametrize("kind,E,H,B,Tq,Td,tag", CONFIG_CASES)
def test_config_sized_parity_vs_port(kind, E, H, B, Tq, Td, tag):
    torch.manual_seed(0)
    port = rp.disable_dropout(rp.build_port(kind, E, H)).train()
    m = _no_dropout(CLS[kind](E, H))
    m.load_state_dict(port.state_dict())
    m.to(DEV).train()
    g = torch.Generator().manual_seed(1234)
    q, d = torch.randn(B, Tq, E, generator=g) * 0.1, torch.randn(B, Td, E, generator=g) * 0.1
    q[1, Tq // 2:] = 0
    d[2, Td // 4:] = 0                                   # zero tail padding is processed
    qv_r, dv_r = port(q, d)
    loss_r = PORT_LOSSES[tag](qv_r, dv_r)
    loss_r.backward()
    qv, dv = m(q.to(DEV), d.to(DEV))
    assert rel_err(qv, qv_r.detach()) < TOL and rel_err(dv, dv_r.detach()) < TOL
    loss = LOSSES[tag]()(qv, dv)
    assert abs(loss.item() - loss_r.item()) <= TOL * abs(loss_r.item())
    loss.backward()
    ref = dict(port.named_parameters())
    scale = max(p.grad.norm().item() for p in ref.values() if p.grad is not None)
    worst = 0.0
    for k, p in m.named_parameters():
        rg = ref[k].grad if ref[k].grad is not None else torch.zeros_like(ref[k])
        assert _grad_close(p.grad, rg, scale), (k, rel_err(p.grad, rg))
        if rg.norm() > 1e-3 * scale:
            worst = max(worst, rel_err(p.grad, rg))
    print(f"[parity] {kind} {tag}: emb {rel_err(qv, qv_r.detach()):.2e}/{rel_err(dv, dv_r.detach()):.2e} "
          f"loss {abs(loss.item() - loss_r.item()) / abs(loss_r.item()):.2e} worst grad {worst:.2e}")


def test_margin_model_mode_switch_and_shared_projection():
    m = tt.TwoTowerModel(20, 8).to(DEV)
    q, d = torch.randn(5, 4, 20, device=DEV), torch.randn(5, 6, 20, device=DEV)
    m.eval()
    sim = m(q, d)
    assert sim.shape == (5, 5) and sim.abs().max() <= 1 + 1e-5
    _no_dropout(m).train()
    qv, dv = m(q, d)
    assert torch.allclose(qv.norm(dim=1), torch.ones(5, device=DEV), atol=1e-5)
    assert rel_err(qv @ dv.t(), sim) < 1e-5
    (qv.sum() + dv.sum()).backward()
    assert m.projection[0].weight.grad is not None


def test_token_id_front_end_matches_float_path_and_token_model():
    torch.manual_seed(0)
    table = torch.randn(100, 20)
    m = tt.EnhancedTwoTowerModel(20, 6).to(DEV).eval().attach_embeddings(table.to(DEV), pad_id=0)
    ids_q, ids_d = torch.randint(0, 100, (7, 5)), torch.randint(0, 100, (7, 9))
    tz = table.clone(); tz[0] = 0
    with torch.no_grad():
        a = m(ids_q.to(DEV), ids_d.to(DEV))
        b = m(tz[ids_q].to(DEV), tz[ids_d].to(DEV))
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    # token model of train_two_tower.py: uni-GRU, final hidden state, triplet API
    tok = tt.TokenTwoTowerModel(vocab_size=100, embedding_dim=20, hidden_dim=12, pretrained_embeddings=table).to(DEV)
    ref = torch.nn.GRU(20, 12, batch_first=True)
    ref.load_state_dict({k: v.cpu() for k, v in tok.doc_encoder.state_dict().items()})
    qe, pe, ne = tok(ids_q.to(DEV), ids_d.to(DEV), ids_d.flip(0).to(DEV))
    assert qe.shape == (7, 12) and ne.shape == (7, 12)
    assert rel_err(pe, ref(tz[ids_d])[1][-1].detach()) < 1e-5
    loss = torch.nn.TripletMarginLoss(margin=0.2, p=2)(qe, pe, ne)
    loss.backward()
    assert tok.query_encoder.weight_hh_l0.grad is not None


def test_lengths_option_matches_pack_padded_sequence():
    torch.manual_seed(2)
    port = rp.build_port("enhanced", 20, 6).eval()
    m = tt.EnhancedTwoTowerModel(20, 6)
    m.load_state_dict(port.state_dict())
    m.to(DEV).eval()
    x = torch.randn(5, 8, 20)
    lens = torch.tensor([8, 3, 1, 5, 8])
    pk = torch.nn.utils.rnn.pack_padded_sequence(x, lens, batch_first=True, enforce_sorted=False)
    with torch.no_grad():
        h_n = port.doc_encoder(pk)[1]
        ref = port.doc_proj(torch.cat((h_n[-2], h_n[-1]), 1))
        got = m.encode_doc(x.to(DEV), lengths=lens)
    assert rel_err(got, ref) < 1e-5


def test_train_step_with_optimizer_and_clip_reduces_loss():
    """The reference loop body (simple_two_tower.py:229-243) runs unchanged on the drop-in."""
    torch.manual_seed(0)
    m = tt.SimpleTwoTower(20, 16).to(DEV)
    crit = S.InfoNCELoss(0.1)
    opt = torch.optim.AdamW(m.parameters(), lr=1e-2, weight_decay=0.01)
    q = torch.randn(32, 6, 20, device=DEV)
    d = q + 0.1 * torch.randn_like(q)
    losses = []
    for _ in range(12):
        opt.zero_grad()
        qv, dv = m(q, d)
        loss = crit(qv, dv)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(m.parameters(), 1.0)
        opt.step()
        losses.append(loss.item())
    assert losses[-1] < 0.7 * losses[0]


def test_retrieval_pipeline_mrr():
    torch.manual_seed(4)
    docs = torch.nn.functional.normalize(torch.randn(5000, 64), dim=1)
    rel = torch.randint(0, 5000, (200,))
    q = torch.nn.functional.normalize(docs[rel] + 0.05 * torch.randn(200, 64), dim=1)
    val, idx = tt.search_topk(q.to(DEV), docs.to(DEV), k=10)
    rv, ri = rp.search_topk(q, docs, 10)
    assert (idx.cpu() == ri).float().mean().item() > 0.999     # distinct random scores; fp32 rounding may swap a near-tie
    out = tt.mrr_recall(idx, rel.to(DEV))
    ref = rp.mrr_recall(ri, rel)
    assert abs(out["MRR"] - ref["MRR"]) < 1e-12 and out["MRR"] > 0.9
