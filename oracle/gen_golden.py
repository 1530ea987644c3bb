"""TEST INFRASTRUCTURE ONLY - generate tests/golden/*.npz by running the UNMODIFIED reference.

Run in the build container (needs /root/reference):  python oracle/gen_golden.py
The reference has no tests / golden vectors of its own, so parity is pinned
by outputs of the reference's nn.Modules / loss classes executed here on CPU fp32 with
seeded inputs.  Small dims keep the fixtures a few hundred KB; the config-sized parity
tests use oracle/ref_port.py, itself pinned by these files.

Each model file holds: state_dict (sd/<key>), inputs q,d, eval-mode encode_query/encode_doc,
eval-mode forward, train-mode forward with dropout p forced to 0, every loss the
reference defines for that model and d(loss)/d(param) for all parameters.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.ref_import import load_reference  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def zero_dropout(m):
    for mod in m.modules():
        if isinstance(mod, torch.nn.Dropout):
            mod.p = 0.0
        if isinstance(mod, torch.nn.GRU):
            mod.dropout = 0.0


def pad_tail(x, lens):
    # dataset contract (simple_two_tower.py:103-108): zero rows appended at the END, still processed
    for b, n in enumerate(lens):
        x[b, n:] = 0
    return x


def model_case(name, cls, E, H, B, Tq, Td, losses, seed):
    torch.manual_seed(seed)
    m = cls(E, H)
    g = torch.Generator().manual_seed(seed + 1)
    q = pad_tail(torch.randn(B, Tq, E, generator=g) * 0.5, [Tq, Tq - 2, 1, Tq // 2] + [Tq] * (B - 4))
    d = pad_tail(torch.randn(B, Td, E, generator=g) * 0.5, [Td, 3, Td - 1, Td] + [Td] * (B - 4))
    rec = {"E": E, "H": H, "q": q.numpy(), "d": d.numpy()}
    for k, v in m.state_dict().items():
        rec["sd/" + k] = v.numpy().copy()
    m.eval()
    with torch.no_grad():
        rec["eval/encode_query"] = m.encode_query(q).numpy()
        rec["eval/encode_doc"] = m.encode_doc(d).numpy()
        out = m(q, d)
        if isinstance(out, tuple):
            rec["eval/forward_q"], rec["eval/forward_d"] = out[0].numpy(), out[1].numpy()
        else:
            rec["eval/forward_sim"] = out.numpy()
    zero_dropout(m)
    m.train()
    for lname, crit in losses.items():
        m.zero_grad()
        qv, dv = m(q, d)
        loss = crit(qv, dv)
        loss.backward()
        rec[f"train/{lname}/q"] = qv.detach().numpy().copy()
        rec[f"train/{lname}/d"] = dv.detach().numpy().copy()
        rec[f"train/{lname}/loss"] = np.float32(loss.item())
        for k, p in m.named_parameters():
            rec[f"train/{lname}/grad/{k}"] = (p.grad if p.grad is not None else torch.zeros_like(p)).numpy().copy()
    np.savez_compressed(os.path.join(OUT, f"{name}.npz"), **rec)
    print(name, "params", sum(p.numel() for p in m.parameters()), "entries", len(rec))


def main():
    os.makedirs(OUT, exist_ok=True)
    S, M, E = load_reference()
    model_case("simple_E300_H8", S.SimpleTwoTower, 300, 8, 6, 7, 9,
               {"infonce_sym_t0.1": S.InfoNCELoss(0.1)}, 11)
    model_case("margin_E300_H8", M.TwoTowerModel, 300, 8, 6, 7, 9,
               {"infonce_q2d_t0.1": M.InfoNCELoss(0.1),
                "margin_ce_m0.2_t0.1": E.MarginRankingLoss(0.2, 0.1)}, 12)
    model_case("enhanced_E300_H6", E.EnhancedTwoTowerModel, 300, 6, 6, 7, 9,
               {"infonce_q2d_t0.07": E.InfoNCELoss(),
                "margin_ce_m0.2_t0.1": E.MarginRankingLoss()}, 13)
    # a second, wider enhanced case with a non-multiple-of-anything hidden size
    model_case("enhanced_E20_H5", E.EnhancedTwoTowerModel, 20, 5, 5, 4, 11,
               {"margin_ce_m0.2_t0.1": E.MarginRankingLoss()}, 14)

    # ---- loss / mining functions on free-standing vectors
    g = torch.Generator().manual_seed(99)
    B, D, k = 9, 16, 3
    q = torch.randn(B, D, generator=g, requires_grad=True)
    p = torch.randn(B, D, generator=g, requires_grad=True)
    neg = torch.randn(B * k, D, generator=g, requires_grad=True)
    rec = {"q": q.detach().numpy(), "pos": p.detach().numpy(), "neg": neg.detach().numpy()}

    def record(tag, loss, wrt):
        grads = torch.autograd.grad(loss, wrt)
        rec[f"{tag}/loss"] = np.float32(loss.item())
        for nme, gr in zip(("dq", "dpos", "dneg"), grads):
            rec[f"{tag}/{nme}"] = gr.numpy().copy()

    record("simple_infonce_t0.1", S.InfoNCELoss(0.1)(torch.nn.functional.normalize(q), torch.nn.functional.normalize(p)), (q, p))
    record("margin_infonce_t0.07", M.InfoNCELoss()(q, p), (q, p))
    record("enhanced_infonce_t0.07", E.InfoNCELoss()(q, p), (q, p))
    record("margin_ce_m0.2_t0.1", E.MarginRankingLoss()(q, p), (q, p))
    record("margin_ce_m0.5_t0.05", E.MarginRankingLoss(0.5, 0.05)(q, p), (q, p))
    record("hinge_m0.2", E.MarginRankingLoss()(q, p, neg), (q, p, neg))
    record("hinge_m1.0", E.MarginRankingLoss(margin=1.0)(q, p, neg), (q, p, neg))
    # hard negatives: N docs, distinct scores so torch.topk is well defined
    N = 40
    docs = torch.randn(N, D, generator=g)
    rec["docs"] = docs.numpy()
    for kk in (5, 8):
        rec[f"hardneg_k{kk}"] = np.stack([
            E.get_hard_negatives(q.detach()[i], docs, positive_idx=i + 4, k=kk).numpy() for i in range(B)])
    rec["hardneg_pos_offset"] = np.int64(4)
    # retrieval scan semantics (simple_two_tower.py:333-340): cosine + topk(10)
    sims = torch.stack([torch.nn.functional.cosine_similarity(q.detach()[i:i + 1], docs) for i in range(B)])
    top = torch.topk(sims, k=10)
    rec["search_top10_idx"], rec["search_top10_val"] = top.indices.numpy(), top.values.numpy()
    np.savez_compressed(os.path.join(OUT, "losses_mining.npz"), **rec)
    print("losses_mining entries", len(rec))


if __name__ == "__main__":
    main()
