"""TEST INFRASTRUCTURE ONLY - explicit numpy restatement of the per-kernel math.

Secondary oracle for kernel-level unit tests (small shapes; float64 by default so
it is the "true" answer both the CUDA kernels and the torch port are compared to).
Equations follow SURVEY.md 3.2 / 8(c), which restate what stock torch does at the
reference call sites:
  nn.GRU            simple_two_tower.py:17-32, margin_two_tower.py:12-28, enhanced_two_tower.py:17-33
  head              simple_two_tower.py:35-49, margin_two_tower.py:30-35, enhanced_two_tower.py:36-48
  F.normalize       simple_two_tower.py:55,61; margin_two_tower.py:39-40
  cross_entropy     simple_two_tower.py:76-77; enhanced_two_tower.py:99-100
  cosine_similarity enhanced_two_tower.py:113-118,125 (eps 1e-8)
"""
import numpy as np


def sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def gru_direction(x, w_ih, w_hh, b_ih, b_hh, reverse=False, lengths=None, dtype=np.float64):
    """x [B,T,In] -> (out [B,T,H], h_last [B,H]).  Gate order r,z,n; b_hn inside r*(.).
    lengths: optional [B]; steps t >= len leave h unchanged and emit zeros (pack_padded semantics)."""
    x = x.astype(dtype); w_ih = w_ih.astype(dtype); w_hh = w_hh.astype(dtype)
    b_ih = b_ih.astype(dtype); b_hh = b_hh.astype(dtype)
    B, T, _ = x.shape
    H = w_hh.shape[1]
    h = np.zeros((B, H), dtype)
    out = np.zeros((B, T, H), dtype)
    steps = range(T - 1, -1, -1) if reverse else range(T)
    for t in steps:
        gi = x[:, t] @ w_ih.T + b_ih
        gh = h @ w_hh.T + b_hh
        r = sigmoid(gi[:, :H] + gh[:, :H])
        z = sigmoid(gi[:, H:2 * H] + gh[:, H:2 * H])
        n = np.tanh(gi[:, 2 * H:] + r * gh[:, 2 * H:])
        h_new = (1.0 - z) * n + z * h
        if lengths is not None:
            live = (t < np.asarray(lengths))[:, None]
            h = np.where(live, h_new, h)
            out[:, t] = np.where(live, h_new, 0.0)
        else:
            h = h_new
            out[:, t] = h
    return out, h


def gru_stack(x, params, prefix, num_layers=2, bidirectional=True, lengths=None, dtype=np.float64):
    """params: dict name->array with nn.GRU names under `prefix.`; returns (out, h_n [L*D,B,H])."""
    inp = x
    h_n = []
    for l in range(num_layers):
        outs = []
        for sfx, rev in (("", False), ("_reverse", True))[: 2 if bidirectional else 1]:
            g = lambda n: params[f"{prefix}.{n}_l{l}{sfx}"]
            o, h = gru_direction(inp, g("weight_ih"), g("weight_hh"), g("bias_ih"), g("bias_hh"),
                                 reverse=rev, lengths=lengths, dtype=dtype)
            outs.append(o); h_n.append(h)
        inp = np.concatenate(outs, axis=2)
    return inp, np.stack(h_n)


def layer_norm(x, w, b, eps=1e-5):
    mu = x.mean(-1, keepdims=True)
    var = ((x - mu) ** 2).mean(-1, keepdims=True)      # biased
    return (x - mu) / np.sqrt(var + eps) * w + b


def l2_normalize(x, eps=1e-12):
    return x / np.maximum(np.sqrt((x * x).sum(-1, keepdims=True)), eps)


def head(v, params, prefix, kind, dtype=np.float64):
    """kind: simple (Lin,LN,ReLU,Lin,+normalise) | margin (Lin,LN,ReLU) | enhanced (Lin,LN,ReLU,Lin)."""
    p = lambda n: params[f"{prefix}.{n}"].astype(dtype)
    y = v.astype(dtype) @ p("0.weight").T + p("0.bias")
    y = np.maximum(layer_norm(y, p("1.weight"), p("1.bias")), 0.0)
    if kind == "simple":
        return l2_normalize(y @ p("4.weight").T + p("4.bias"))
    if kind == "enhanced":
        return y @ p("3.weight").T + p("3.bias")
    return y


def tower(x, params, enc_prefix, head_prefix, kind, lengths=None, dtype=np.float64):
    _, h_n = gru_stack(x, params, enc_prefix, lengths=lengths, dtype=dtype)
    return head(np.concatenate((h_n[-2], h_n[-1]), axis=1), params, head_prefix, kind, dtype)


def cross_entropy_diag(logits, pos_offset=0):
    m = logits.max(1, keepdims=True)
    lse = np.log(np.exp(logits - m).sum(1)) + m[:, 0]
    idx = np.arange(logits.shape[0])
    return (lse - logits[idx, idx + pos_offset]).mean()


def score_ce(x, y, inv_temp, margin=0.0, pos_offset=0, normalize=False):
    """Generic in-batch CE primitive the CUDA score kernel implements:
    logits[i,j] = <x_i,y_j>*inv_temp - margin*(j != pos_offset+i); mean_i CE(logits_i, pos_offset+i)."""
    if normalize:
        x, y = l2_normalize(x), l2_normalize(y)
    s = (x @ y.T) * inv_temp
    mask = np.ones_like(s); mask[np.arange(x.shape[0]), np.arange(x.shape[0]) + pos_offset] = 0.0
    return cross_entropy_diag(s - margin * mask, pos_offset)


def cosine_rows(q, docs, eps=1e-8):
    qn = np.maximum(np.sqrt((q * q).sum(-1)), eps)
    dn = np.maximum(np.sqrt((docs * docs).sum(-1)), eps)
    return (q @ docs.T) / (qn[:, None] * dn[None, :])


def stable_topk(scores, k):
    order = np.argsort(-scores, axis=-1, kind="stable")[..., :k]
    return np.take_along_axis(scores, order, -1), order


def hard_negatives_rows(q, docs, pos_offset=0, k=5):
    sim = cosine_rows(q, docs)
    sim[np.arange(q.shape[0]), np.arange(q.shape[0]) + pos_offset] = -1.0
    return stable_topk(sim, k)[1]


def hinge_explicit_negatives(q, pos, neg, margin=0.2, eps=1e-8):
    B, D = q.shape
    k = neg.shape[0] // B
    cos = lambda a, b: (a * b).sum(-1) / (np.maximum(np.sqrt((a * a).sum(-1)), eps) *
                                           np.maximum(np.sqrt((b * b).sum(-1)), eps))
    pos_sim = cos(q, pos)
    neg_sim = cos(q[:, None, :], neg.reshape(B, k, D)).mean(1)
    return np.maximum(margin - pos_sim + neg_sim, 0.0).mean()
