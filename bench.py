#!/usr/bin/env python
"""bench.py - train query-doc pairs/s (fwd + loss + hard-negative mining + bwd + optimizer) of the two_towers hot path.

  python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (N>1: launched by torchrun)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port) on the host cores

Workload (BASELINE.json configs[2] at N=1, configs[3] at N>1, weak scaling): EnhancedTwoTowerModel(300, 128)
(bi-GRU hidden 256, 2 layers, LayerNorm projection head), MarginRankingLoss in-batch branch, row top-8 hard
negatives, B = 4096 pairs per GPU, query length 32, doc length 128, synthetic token ids over a 3 000 000 x 300
frozen table (the size of word2vec-google-news-300), random-init weights.  One JSON line on stdout (rank 0).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

E, HID, TQ, TD, TOPK = 300, 128, 32, 128, 8
GRU_H = 2 * HID
METRIC = "train query-doc pairs/sec (fwd+bwd)"
# algorithmic GEMM flops per pair, fwd+bwd = 3 x fwd: 12*Hg*(300+4*Hg) per token per tower
FLOPS_PER_PAIR = 3.0 * 12 * GRU_H * (E + 4 * GRU_H) * (TQ + TD)


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="ours", choices=["ours", "reference", "torch-gpu"])
    p.add_argument("--batch", type=int, default=4096, help="pairs per GPU (4096 = the BASELINE config)")
    p.add_argument("--vocab", type=int, default=3_000_000)
    p.add_argument("--cpu-batch", type=int, default=256, help="pairs per step of the bounded CPU sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--optimizer", default="fused", choices=["fused", "torch"],
                   help="fused = this repo's tt_opt_step Adam kernel (FusedAdam); torch = torch.optim.Adam(fused=True)")
    p.add_argument("--serial-towers", action="store_true", help="both towers on one stream (for per-kernel launch lists)")
    p.add_argument("--eval-queries", type=int, default=100_000, help="queries of the retrieval leg (C5: 100 000; 0 = skip it)")
    p.add_argument("--no-graph", action="store_true", help="eager steps instead of the CUDA-graph captured step")
    p.add_argument("--no-configs", action="store_true", help="skip the secondary configurations (C1', C2, C3b)")
    p.add_argument("--no-torch-gpu", action="store_true", help="skip the stock-torch-on-the-GPU arm folded into the line")
    p.add_argument("--eval-docs", type=int, default=1_100_000, help="corpus rows per GPU of the retrieval leg (C5 shard)")
    return p.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return dict(hbm=d["hbm_gbs"], bf16_burst=d["bf16_tflops"], bf16_sustained=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, bf16_burst=1590.0, bf16_sustained=1400.0, src="fallback")


def make_pair_ids(B, Tq, Td, V, gen):
    """A (query, relevant doc) batch: the doc is random text that CONTAINS the query's tokens (a contiguous window at a random
    position), as a relevant passage shares words with its query.  With independent random queries and docs there is nothing to
    learn: the loss stays at ln(B_global), the embeddings drift together and the in-batch scores become near-ties, which is
    neither what training looks like nor neutral for the timing (rows whose top-k cannot be separated within the tensor
    route's error bound take the exact re-scan)."""
    q, d = make_ids(B, Tq, V, gen), make_ids(B, Td, V, gen)
    qlen, dlen = (q != 0).sum(1), (d != 0).sum(1)
    n = torch.minimum(qlen, dlen)
    start = (torch.rand(B, generator=gen) * (dlen - n + 1).float()).long()
    j = torch.arange(Tq).unsqueeze(0)
    pos = (start.unsqueeze(1) + j).clamp_(max=Td - 1)
    keep = j < n.unsqueeze(1)
    d.scatter_(1, pos, torch.where(keep, q, d.gather(1, pos)))
    return q, d


def make_ids(B, T, V, gen):
    """Power-law (Zipf-like) ids in [1,V), length ~U[T/4,T], pad id 0 after the length (still processed)."""
    u = torch.rand(B, T, generator=gen)
    ids = torch.exp(u * torch.log(torch.tensor(float(V)))).long().clamp_(1, V - 1)
    lens = torch.randint(max(1, T // 4), T + 1, (B, 1), generator=gen)
    ids[torch.arange(T).unsqueeze(0) >= lens] = 0
    return ids


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index
        self.t0 = self.t1 = None

    def start(self):
        """Launch the sampler (nvidia-smi needs ~0.2 s before its first line, so this happens before the warm-up);
        only lines that arrive inside [mark_begin, mark_end] are kept."""
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([time.perf_counter()] + [c.strip() for c in line.split(",")])

    def mark_begin(self):
        self.t0 = time.perf_counter()

    def mark_end(self):
        self.t1 = time.perf_counter()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        if self.t1 is None:
            self.mark_end()
        time.sleep(0.05)
        self.proc.terminate()
        rows = [r[1:] for r in self.rows if self.t0 is None or self.t0 <= r[0] <= self.t1 + 0.03]
        sm = [float(r[1]) for r in rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows if len(r) >= 8 for n, v in zip(names, r[4:8]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ----------------------------------------------------------------------------------------- CPU reference arm
def cpu_reference(steps, warmup, B, seed=1234):
    """The reference's own CPU implementation of the path: its nn.Modules restated over stock torch CPU ops
    (oracle/ref_port.py, pinned to the unmodified reference by tests/golden), fp32, train() mode, all host
    threads; one step = fwd + MarginRankingLoss + bwd + Adam step on a bounded sample of B pairs
    (pre-embedded float inputs, as the reference's Dataset hands them over).  Returns (pairs/s, ms/step, threads)."""
    from oracle import ref_port as rp
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    model = rp.build_port("enhanced", E, HID).train()
    opt = torch.optim.Adam(model.parameters())
    gen = torch.Generator().manual_seed(seed)
    V = 13085
    table = torch.randn(V, E, generator=gen) * 0.1
    table[0] = 0
    batches = [tuple(table[t] for t in make_pair_ids(B, TQ, TD, V, gen)) for _ in range(2)]
    times = []
    for i in range(warmup + steps):
        q, d = batches[i % 2]
        t0 = time.perf_counter()
        opt.zero_grad()
        qv, dv = model(q, d)
        loss = rp.margin_ce_inbatch(qv, dv, 0.2, 0.1)
        with torch.no_grad():
            rp.stable_topk(torch.nn.functional.cosine_similarity(qv.unsqueeze(1), dv.unsqueeze(0), dim=2)
                           .masked_fill(torch.eye(B, dtype=torch.bool), -1.0), TOPK)
        loss.backward()
        opt.step()
        loss.item()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    ms = 1e3 * sum(times) / len(times)
    return B / (ms / 1e3), ms, threads


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    val, ms, threads = cpu_reference(args.steps, args.warmup, args.cpu_batch)
    sample = (f"{args.cpu_batch} pairs/step (BASELINE batch is {args.batch}/GPU), Tq/Td {TQ}/{TD}, pre-embedded fp32 "
              f"inputs, {args.steps} timed steps after {args.warmup} warm-up")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": dict(workload_config(args.batch, args.gpus, args.vocab), cpu_sample_batch=args.cpu_batch),
        "cpu_baseline": {"value": val, "unit": "pairs/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------- stock torch on the GPU
def torch_gpu_arm(args, steps, warmup):
    """Extra arm (SURVEY 8d "the real beat-it bar"), not part of the driver contract: the reference's architecture
    (enhanced_two_tower.py:14-65: nn.GRU bi x2 + Linear/LayerNorm/ReLU/Linear heads), its in-batch margin loss
    (:90-101) and cosine + topk mining (:123-133) written with STOCK torch modules and run on cuda:0 - cuDNN GRU,
    cuBLAS - on the same workload as the default arm (ids gathered from the same-sized table on the device).
    Nothing of this repo's library and nothing under oracle/ is on this path.  One JSON line per cuDNN TF32 setting."""
    import torch.nn as nn
    import torch.nn.functional as F
    dev = torch.device("cuda", torch.cuda.current_device())
    rows = []
    B, V = args.batch, args.vocab

    class Stock(nn.Module):
        def __init__(self):
            super().__init__()
            mk = lambda: nn.GRU(E, GRU_H, num_layers=2, batch_first=True, bidirectional=True, dropout=0.1)
            hd = lambda: nn.Sequential(nn.Linear(4 * HID, 2 * HID), nn.LayerNorm(2 * HID), nn.ReLU(), nn.Linear(2 * HID, HID))
            self.query_encoder, self.doc_encoder, self.query_proj, self.doc_proj = mk(), mk(), hd(), hd()

        def enc(self, x, gru, proj):
            _, h = gru(x)
            return proj(torch.cat([h[-2], h[-1]], dim=1))

        def forward(self, q, d):
            return self.enc(q, self.query_encoder, self.query_proj), self.enc(d, self.doc_encoder, self.doc_proj)

    gen = torch.Generator().manual_seed(1234)
    table = torch.empty(V, E, device=dev).normal_(0, 0.1, generator=torch.Generator(device=dev).manual_seed(7))
    table[0] = 0
    ids = [tuple(t.to(dev) for t in make_pair_ids(B, TQ, TD, V, gen)) for _ in range(4)]
    eye = torch.eye(B, device=dev)
    for tf32 in (True, False):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.manual_seed(0)
        model = Stock().to(dev).train()
        opt = torch.optim.Adam(model.parameters(), fused=True)

        def step(qi, di):
            opt.zero_grad(set_to_none=True)
            qv, dv = model(F.embedding(qi, table), F.embedding(di, table))
            sim = qv @ dv.t() / 0.1
            sim = sim - 0.2 * (1 - eye)
            loss = F.cross_entropy(sim, torch.arange(B, device=dev))
            with torch.no_grad():
                cs = F.normalize(qv, dim=1, eps=1e-8) @ F.normalize(dv, dim=1, eps=1e-8).t()
                cs.fill_diagonal_(-1.0)
                hard = cs.topk(TOPK, dim=1).indices
            loss.backward()
            opt.step()
            return loss, hard

        for i in range(warmup):
            step(*ids[i % 4])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for i in range(steps):
            loss, _ = step(*ids[i % 4])
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        rows.append({"impl": "torch-gpu", "metric": METRIC, "value": B / (ms * 1e-3), "unit": "pairs/s", "n_gpus": 1,
                     "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True,
                     "dtype": "f32 (cuDNN GRU, allow_tf32=%s; cuBLAS matmul fp32)" % tf32, "data": "synthetic",
                     "config": workload_config(B, 1, V), "loss": loss.item(),
                     "note": "stock torch nn.GRU/nn.Linear/LayerNorm/cross_entropy/topk on the GPU; not this repo's kernels"})
        del model, opt
    torch.backends.cudnn.allow_tf32 = True
    del table, ids, eye
    torch.cuda.empty_cache()
    return rows


def run_torch_gpu(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    torch.cuda.set_device(0)
    for row in torch_gpu_arm(args, args.steps, args.warmup):
        print(json.dumps(row), flush=True)


# ----------------------------------------------------------------------------------------- this repo's arm
def seq_kernel_work(B):
    """Algorithmic work of the persistent recurrence kernels per train step (both towers, 2 layers, 2 directions);
    DESIGN.md section 4.  Per (token, direction, layer), Hg = 256:
      forward  gru_seq_kernel     reads gi fp32 [3Hg] (12 Hg B); writes the five 16-bit gate-record planes z, r, A, C, B
                                  (10 Hg B) and h_t as fp16 hi+lo planes (4 Hg B)                    -> 26 Hg B
      backward gru_bwd_seq_kernel reads the gate records (10 Hg B) [+ dL/dout fp32 4 Hg B, counted];
                                  writes the dG fp16 planes dr|dz|dn*r|dn (8 Hg B)                  -> 22 Hg B
      flops (either)              one [1 x Hg] x [Hg x 3Hg] recurrent product = 6 Hg^2 (one fp32-equivalent pass;
                                  the forward issues 3 fp16 tensor passes of it, the backward 1)."""
    units = B * (TQ + TD) * 2 * 2
    return {0: dict(bytes=units * 26 * GRU_H, flops=units * 6 * GRU_H * GRU_H, passes=3),
            1: dict(bytes=units * 22 * GRU_H, flops=units * 6 * GRU_H * GRU_H, passes=1)}


def kernel_rooflines(lib, steps, B, pk):
    """Per-family device time from CUDA events the library recorded around every launch, on the launch stream,
    inside the timed region (tt_prof_*).  The dominant family becomes the `roofline` object."""
    import ctypes
    work = seq_kernel_work(B)
    fams = []
    for f in range(3):
        ms, n = ctypes.c_double(0), ctypes.c_uint64(0)
        lib.tt_prof_read(f, ctypes.byref(ms), ctypes.byref(n))
        fams.append(dict(family=f, kernel=lib.tt_prof_family(f).decode(), ms_per_step=ms.value / steps,
                         launches_per_step=n.value / steps))
    traffic = {}
    tpath = os.path.join(ROOT, "profiles", "dram_traffic.json")
    if os.path.exists(tpath):
        traffic = json.load(open(tpath))
    out = []
    for fam in fams[:2]:
        w = work[fam["family"]]
        sec = fam["ms_per_step"] * 1e-3
        if sec <= 0:
            continue
        nl = max(fam["launches_per_step"], 1)
        gbs = w["bytes"] / sec / 1e9
        tf = w["flops"] / sec / 1e12
        t = traffic.get(str(fam["family"]))
        out.append({"bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s", "frac": gbs / pk["hbm"],
                    "traffic": (t["dram_bytes_per_step"] / nl) if t else None,
                    "algorithmic_bytes_per_launch": w["bytes"] / nl,
                    "kernel": fam["kernel"], "ms_per_launch": fam["ms_per_step"] / nl,
                    "launches_per_step": fam["launches_per_step"], "ms_per_step": fam["ms_per_step"],
                    "tensor": {"algorithmic_tflops": tf, "issued_tflops": tf * w["passes"],
                               "frac_of_bf16_sustained": tf / pk["bf16_sustained"],
                               "issued_frac_of_bf16_sustained": tf * w["passes"] / pk["bf16_sustained"]},
                    "peak_source": f"HBM copy bandwidth, {pk['src']} (MEASURED_PEAKS.json); kernel timed inside the step",
                    "timed": "CUDA events around every launch on its stream, over K steps run with both towers on one stream "
                             "right after the timed region (the headline steps overlap the towers on two streams)",
                    "why_hbm": "at the saved-gate formats this kernel's HBM floor exceeds its tensor floor (DESIGN.md section 4)"})
    out.sort(key=lambda r: -r["ms_per_step"])
    return out, fams


def gemm_table(lib, steps, pk):
    """Per-shape rows of the tcgen05 GEMM engine (family 2) from the same profiled steps: the roofline of the largest family.
    Flops of a row = 2 M N K x batched problems (algorithmic) and x passes (issued); peak = sustained bf16 (fp16 runs at the
    same rate)."""
    from two_towers_b200 import _lib as L
    buf = (L.ProfGemmRow * 64)()
    n = lib.tt_prof_gemm_rows(buf, 64)
    rows = []
    for r in buf[:n]:
        if r.launches == 0 or r.total_ms <= 0:
            continue
        sec = r.total_ms * 1e-3
        fl = 2.0 * r.M * r.N * r.K * r.nbatch * r.launches
        kind = {0: "row-major", 1: "tiled", 2: "top-k selection"}.get((r.flags >> 4) & 15, "?")
        rows.append({"M": r.M, "N": r.N, "K": r.K, "batched": r.nbatch, "passes": r.npass,
                     "layout": ("A^T " if r.flags & 1 else "A ") + ("B MN-major" if r.flags & 2 else "B K-major"), "epilogue": kind,
                     "split_k": bool(r.flags & 256), "launches_per_step": r.launches / steps, "ms_per_step": r.total_ms / steps,
                     "algorithmic_tflops": fl / sec / 1e12, "issued_tflops": fl * r.npass / sec / 1e12,
                     "frac_of_bf16_sustained": fl / sec / 1e12 / pk["bf16_sustained"],
                     "issued_frac_of_bf16_sustained": fl * r.npass / sec / 1e12 / pk["bf16_sustained"]})
    rows.sort(key=lambda x: -x["ms_per_step"])
    tot_ms = sum(x["ms_per_step"] for x in rows)
    tot_fl = sum(x["algorithmic_tflops"] * x["ms_per_step"] for x in rows)
    tot_is = sum(x["issued_tflops"] * x["ms_per_step"] for x in rows)
    summary = None
    if tot_ms > 0:
        summary = {"bound": "tensor", "kernel": "umma_gemm_kernel (tcgen05 GEMM engine), all shapes of the step",
                   "achieved": tot_is / tot_ms, "peak": pk["bf16_sustained"], "unit": "TFLOP/s", "frac": tot_is / tot_ms / pk["bf16_sustained"],
                   "algorithmic_tflops": tot_fl / tot_ms, "algorithmic_frac": tot_fl / tot_ms / pk["bf16_sustained"],
                   "ms_per_step": tot_ms, "peak_source": f"cuBLAS bf16 sustained, {pk['src']} (MEASURED_PEAKS.json)",
                   "note": "achieved counts the tensor passes issued (3 per forward product: fp16 hi/lo split for fp32-grade results)"}
    return rows, summary


def eval_cpu_baseline(q, docs, nq=1000):
    """SURVEY 8d CPU figure for C5: the reference's retrieval semantics (validate_enhanced.py:75-76: cosine_similarity of a
    query against every doc embedding + topk(10)) batched as q_tile @ docs.T on all host threads, nq queries x one shard."""
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    qc, dc = q[:nq].float().cpu(), docs.float().cpu()
    t0 = time.perf_counter()
    dn = torch.nn.functional.normalize(dc, dim=1, eps=1e-8)
    t_norm = time.perf_counter() - t0
    t0 = time.perf_counter()
    out = []
    for i in range(0, nq, 250):
        sim = torch.nn.functional.normalize(qc[i:i + 250], dim=1, eps=1e-8) @ dn.t()
        out.append(sim.topk(10, dim=1).indices)
    sec = time.perf_counter() - t0
    return {"value": nq / sec, "unit": "queries/s", "cores": threads, "kind": "port",
            "sample": f"{nq} queries x {dc.shape[0]} docs (one C5 shard) x {dc.shape[1]}: cosine + topk(10) with stock torch on the "
                      f"host ({sec * 1e3:.0f} ms; corpus normalisation {t_norm * 1e3:.0f} ms outside, as the reference encodes once)"}, \
        torch.cat(out)


def eval_leg(args, dev, rank, world, group):
    """SURVEY 8d metric (ii), BASELINE configs[4]: score + top-10 + shard merge + MRR@10 over pre-computed embeddings.
    Q = 100 000 replicated queries (processed in tiles of 8192), corpus sharded by contiguous global id with 1.1 M rows per
    GPU (the C5 shard size: 8.8 M docs on 8 GPUs); one relevant doc per query planted as normalize(q + 0.5 randn)."""
    import two_towers_b200 as tt
    Q, Nl, D = args.eval_queries, args.eval_docs, HID
    QT = 8192
    g = torch.Generator(device=dev).manual_seed(4321)
    q = torch.nn.functional.normalize(torch.randn(Q, D, device=dev, generator=g), dim=1)
    gl = torch.Generator(device=dev).manual_seed(99 + rank)
    docs = torch.nn.functional.normalize(torch.randn(Nl, D, device=dev, generator=gl), dim=1)
    rel = torch.randint(0, Nl * world, (Q,), device=dev, generator=g)
    mine = (rel // Nl) == rank
    noisy = torch.nn.functional.normalize(q + 0.5 * torch.randn(Q, D, device=dev, generator=g) / D ** 0.5, dim=1)
    docs[(rel % Nl)[mine]] = noisy[mine]

    # the corpus index (row norms + hi/lo planes of the shard) is built once, like the reference's doc_encodings
    # (simple_two_tower.py:318-331, app.py:38-60); its one-off cost is reported beside the per-scan time
    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    index = tt.CorpusIndex(docs, cosine=True, index_base=rank * Nl)       # warm-up of the build kernels
    torch.cuda.synchronize()
    b0.record()
    index = tt.CorpusIndex(docs, cosine=True, index_base=rank * Nl)
    b1.record()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

    def run(timed=False):
        vals, idxs = [], []
        if timed:
            ev[0].record()
        for i in range(0, Q, QT):                      # local shard scan, query tile by query tile
            v, ix = tt.search_topk(q[i:i + QT], index, k=10)
            vals.append(v)
            idxs.append(ix)
        val, idx = torch.cat(vals), torch.cat(idxs)
        if timed:
            ev[1].record()
        val, idx = tt.merge_shard_topk(val, idx, group)     # all-gather of the per-shard lists + merge (identity at N = 1)
        if timed:
            ev[2].record()
        m = tt.mrr_recall(idx, rel)
        if timed:
            ev[3].record()
        return m, idx

    run()
    torch.cuda.synchronize()
    m, idx = run(timed=True)
    torch.cuda.synchronize()
    parts = torch.tensor([ev[0].elapsed_time(ev[3]), ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])],
                         device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(parts, op=dist.ReduceOp.MAX)
    ms, ms_scan, ms_merge, ms_mrr = parts.tolist()
    pk = peaks()
    tf = 2.0 * Q * Nl * D / (ms_scan * 1e-3) / 1e12            # per GPU: every rank scans its own shard
    out = {"metric": "eval queries/sec (score + top-10 + shard merge + MRR@10)", "value": Q / (ms * 1e-3), "unit": "queries/s",
           "queries": Q, "docs_per_gpu": Nl, "docs_total": Nl * world, "dim": D, "ms": ms,
           "ms_scan": ms_scan, "ms_allgather_merge": ms_merge, "ms_mrr": ms_mrr, "query_tile": QT,
           "mrr10": m["MRR"], "recall10": m["R@10"], "index_build_ms": b0.elapsed_time(b1),
           "roofline": {"bound": "tensor", "kernel": "umma_gemm_kernel selection epilogue (fused score + top-k candidates)",
                        "achieved": tf * 2, "peak": pk["bf16_sustained"], "unit": "TFLOP/s", "frac": tf * 2 / pk["bf16_sustained"],
                        "algorithmic_tflops": tf, "algorithmic_frac": tf / pk["bf16_sustained"], "passes": 2,
                        "peak_source": f"cuBLAS bf16 sustained, {pk['src']}; the scan time includes the exact fp32 re-score of the "
                                       "candidates and the list merge (the GEMM kernel alone is faster)"},
           "note": "BASELINE configs[4] shape per GPU: 100 000 queries against a 1.1 M-doc shard (8.8 M docs on 8 GPUs); the corpus "
                   "index (norms + fp16 hi/lo planes of the shard) is built once outside the timed scan (index_build_ms), as the "
                   "reference encodes its corpus once"}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cb, cpu_idx = eval_cpu_baseline(q, docs)
        out["cpu_baseline"] = cb
        out["cpu_sample_index_agreement"] = (idx[:cpu_idx.shape[0]].cpu() == cpu_idx).float().mean().item()
    return out


def secondary_configs(args, dev, table, graphed):
    """The other train configurations of BASELINE.json / SURVEY section 8 as short measured lines beside the C3 headline
    (same timing rules: resident ids, CUDA events, warm-up, 4 rotating id batches; whole step incl. the optimizer):
      C1'  SimpleTwoTower(300,128), B = 256, Tq/Td 32/32, symmetric InfoNCE(0.1), AdamW + clip 1.0 (simple_two_tower.py:175-243)
      C2   TwoTowerModel(300,256), B = 1024, Tq/Td 32/128, MarginRankingLoss in-batch and InfoNCELoss(0.1), Adam + clip
           (margin_two_tower.py:9-85, train_margin.py:62-67,111-123)
      C3b  EnhancedTwoTowerModel(300,512) - the reference default hidden_dim (train_enhanced.py:30) - B = 4096, Tq/Td 32/128"""
    import two_towers_b200 as tt
    from two_towers_b200 import simple_two_tower as S
    pk = peaks()
    small = torch.empty(13085, E, device=dev).normal_(0, 0.1, generator=torch.Generator(device=dev).manual_seed(8))
    small[0] = 0
    cases = [
        ("c1_simple", lambda: tt.SimpleTwoTower(E, 128), lambda: S.InfoNCELoss(0.1), 256, 32, 32, 128, small,
         lambda ps, c: tt.FusedAdamW(ps, lr=1e-4, weight_decay=0.01, max_grad_norm=1.0, capturable=c), 20, 0),
        ("c2_margin_ce", lambda: tt.TwoTowerModel(E, 256), lambda: tt.MarginRankingLoss(0.2, 0.1), 1024, 32, 128, 256, table,
         lambda ps, c: tt.FusedAdam(ps, lr=1e-4, max_grad_norm=1.0, capturable=c), 10, 0),
        ("c2_infonce", lambda: tt.TwoTowerModel(E, 256), lambda: tt.InfoNCELoss(0.1), 1024, 32, 128, 256, table,
         lambda ps, c: tt.FusedAdam(ps, lr=1e-4, max_grad_norm=1.0, capturable=c), 10, 0),
        ("c3b_enhanced_h512", lambda: tt.EnhancedTwoTowerModel(E, 512), lambda: tt.MarginRankingLoss(0.2, 0.1), 4096, 32, 128, 1024, table,
         lambda ps, c: tt.FusedAdam(ps, capturable=c), 3, TOPK),
    ]
    out = []
    for name, mk, mkloss, B, Tq, Td, Hg, tab, mkopt, steps, mine_k in cases:
        row = {"name": name, "batch": B, "query_len": Tq, "doc_len": Td, "gru_hidden": Hg, "steps": steps}
        try:
            torch.manual_seed(0)
            model = mk().to(dev).train()
            model.attach_embeddings(tab, pad_id=0)
            crit = mkloss()
            gen = torch.Generator().manual_seed(4321)
            V = tab.shape[0]
            ids = [tuple(t.to(dev) for t in make_pair_ids(B, Tq, Td, V, gen)) for _ in range(4)]
            results = {}
            for mode in (["graph", "eager"] if graphed else ["eager"]):
                opt = mkopt(model.parameters(), mode == "graph")
                if mode == "graph":
                    step = tt.GraphedTrainStep(model, crit, opt, mine_k=mine_k)
                else:
                    def step(q, d, opt=opt):
                        opt.zero_grad(set_to_none=True)
                        qv, dv = model(q, d)
                        loss = crit(qv, dv)
                        loss.backward()
                        opt.step()
                        if mine_k:
                            tt.get_hard_negatives_batch(qv.detach(), dv.detach(), k=mine_k)
                        return loss, None
                for i in range(3):
                    step(*ids[i % 4])
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                e0.record()
                th = time.perf_counter()
                for i in range(steps):
                    loss, _ = step(*ids[i % 4])
                host_ms = (time.perf_counter() - th) * 1e3 / steps
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / steps
                results[mode] = {"ms_per_step": ms, "pairs_per_s": B / (ms * 1e-3), "host_queue_ms_per_step": host_ms,
                                 "loss": float(loss.detach())}
                if mode == "graph":
                    results[mode]["launches_per_replay"] = step.launches_per_replay(*ids[0])
                del opt, step
            best = min(results.values(), key=lambda r: r["ms_per_step"])
            flops = 3.0 * 12 * Hg * (E + 4 * Hg) * (Tq + Td)
            row.update({"value": best["pairs_per_s"], "unit": "pairs/s", "ms_per_step": best["ms_per_step"], "modes": results,
                        "gflop_per_pair": flops / 1e9,
                        "step_frac_of_bf16_sustained": flops * best["pairs_per_s"] / 1e12 / pk["bf16_sustained"]})
            del model, crit, ids
        except Exception as exc:                                   # e.g. out of memory for C3b: reported, not fatal
            row["error"] = repr(exc)[:300]
        torch.cuda.empty_cache()
        out.append(row)
    return out


def run_ours(args):
    import torch.distributed as dist
    import two_towers_b200 as tt
    from two_towers_b200 import _lib as L
    from two_towers_b200 import distributed as dops

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device: the product path has no CPU fallback"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD
    lib = L.lib()
    B, V = args.batch, args.vocab
    torch.manual_seed(0)                                 # same weights on every rank
    model = tt.EnhancedTwoTowerModel(E, HID).to(dev).train()
    gen = torch.Generator().manual_seed(1234 + rank)
    table = torch.empty(V, E, device=dev).normal_(0, 0.1, generator=torch.Generator(device=dev).manual_seed(7))
    table[0] = 0
    model.attach_embeddings(table, pad_id=0)
    model.overlap_towers = not args.serial_towers
    lib.tt_debug_wgrad_overlap(0 if args.serial_towers else 1)
    if os.environ.get("TT_BWD_SEQ_MODE"):
        lib.tt_debug_bwd_seq_mode(int(os.environ["TT_BWD_SEQ_MODE"]))
    if os.environ.get("TT_BWD_SEGMENTS"):
        lib.tt_debug_bwd_segments(int(os.environ["TT_BWD_SEGMENTS"]))
    crit = tt.MarginRankingLoss(margin=0.2, temperature=0.1, group=group)
    graphed = (not args.no_graph) and args.optimizer == "fused"
    opt = tt.FusedAdam(model.parameters(), capturable=graphed) if args.optimizer == "fused" else \
        torch.optim.Adam(model.parameters(), fused=True)
    reducer = dops.GradientAllReducer(None, group, named_params=model.named_parameters()) if world > 1 else None
    nb = 4
    host = [tuple(t.pin_memory() for t in make_pair_ids(B, TQ, TD, V, gen)) for _ in range(nb)]
    resident = [(q.to(dev), d.to(dev)) for q, d in host]
    resident_shape = host[0]

    def eager_step(q, d):
        opt.zero_grad(set_to_none=True)
        qv, dv = model(q, d)
        loss = crit(qv, dv)
        d_all = dops.all_gather_rows(dv.detach(), group) if world > 1 else dv.detach()
        hard = tt.get_hard_negatives_batch(qv.detach(), d_all, k=TOPK, pos_offset=rank * B)
        loss.backward()
        if reducer is not None:
            reducer.finish()
        opt.step()
        return loss, hard

    # N = 1: the whole step (zero_grad, both towers, loss, backward, clip-free Adam, hard-negative mining) as ONE CUDA graph,
    # captured on the first call, replayed afterwards (two_towers_b200/graph.py GraphedTrainStep).
    # N > 1: only the towers are graphed (GraphedTowers: forward and backward of both towers as two graphs); the loss with its
    # all-gather / reduce-scatter, the gradient all-reduce and the optimizer stay eager - NCCL calls captured inside a graph
    # cost ~2.5 ms per step at 8 GPUs (profiles/r2_summary.md).  --no-graph keeps the fully eager loop.
    gstep = towers = None
    if graphed and world == 1:
        gstep = tt.GraphedTrainStep(model, crit, opt, reducer=reducer, mine_k=TOPK, pos_offset=rank * B)
    elif graphed:
        towers = tt.GraphedTowers(model)

    seg = []                                  # TT_E2E_DEBUG=1: CUDA events at the phase boundaries of every step
    dbg_losses = []

    def _mark(row):
        if os.environ.get("TT_E2E_DEBUG") == "1":
            e = torch.cuda.Event(enable_timing=True)
            e.record(torch.cuda.current_stream())
            row.append(e)

    def towers_step(q, d):
        row = []
        _mark(row)
        opt.zero_grad(set_to_none=True)
        qv, dv = towers(q, d)
        _mark(row)
        loss = crit(qv, dv)
        _mark(row)
        docs_all = dops.all_gather_rows(dv.detach(), group)
        _mark(row)
        hard = tt.get_hard_negatives_batch(qv.detach(), docs_all, k=TOPK, pos_offset=rank * B)
        _mark(row)
        loss.backward()
        _mark(row)
        reducer.finish()
        _mark(row)
        opt.step()
        _mark(row)
        seg.append(row)
        if os.environ.get("TT_E2E_DEBUG") == "1":
            dbg_losses.append(loss.detach().clone())
        return loss, hard

    def seg_report(tag, rows):
        if not rows or not rows[0] or rank != 0:
            return
        torch.cuda.synchronize()
        names = ["towers fwd (graph)", "loss fwd (all-gather, CE)", "all-gather of the docs", "mining top-k", "backward (CE, reduce-scatter, towers graph)",
                 "reducer.finish", "optimizer"]
        n = len(rows)
        avg = [sum(r[j].elapsed_time(r[j + 1]) for r in rows) / n for j in range(len(names))]
        gaps = sum(a[-1].elapsed_time(b[0]) for a, b in zip(rows, rows[1:])) / max(1, n - 1)
        print(f"[segments] {tag}: " + "; ".join(f"{nm} {v:.2f}" for nm, v in zip(names, avg)) + f"; between steps {gaps:.2f} ms", file=sys.stderr)

    step = gstep if gstep is not None else (towers_step if towers is not None else eager_step)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # Both timed loops run the SAME training steps (W warm-up + K timed from the same initial weights and optimizer state): the
    # step is not data-independent - on synthetic pairs this recipe collapses to constant scores after a few dozen steps
    # (loss = ln(B_global) - margin), every row's top-8 is then a near-tie below the tensor route's error bound and the mining
    # takes the exact scan for ALL rows (0.9 -> 4.2 ms at 8 GPUs, tools/mining_probe.py) - so a loop that starts where the
    # other one ended, or that draws other dropout masks, would measure a different mix of the two regimes.
    init_params = [p.detach().clone() for p in model.parameters()]

    def reset_training_state():
        with torch.no_grad():
            for p, v in zip(model.parameters(), init_params):
                p.copy_(v)                                  # in place: captured graphs, buckets and Adam tables keep their pointers
            for stt in opt.state.values():
                for v in stt.values():
                    if torch.is_tensor(v):
                        v.zero_()
            for obj in (gstep, towers):                     # same dropout masks again: the device-side seed offset of the graphs
                if obj is not None and getattr(obj, "_seed", None) is not None:
                    obj._seed.zero_()

    sampler = ClockSampler(local)
    sampler.start()
    for i in range(args.warmup):
        step(*resident[i % nb])
    # ---- value: inputs resident in HBM, CUDA events on the launch stream, max over ranks
    st = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.mark_begin()
    k0 = lib.tt_kernel_launches()
    e0.record(st)
    th0 = time.perf_counter()
    for i in range(args.steps):
        loss, _ = step(*resident[i % nb])
    host_launch_ms = (time.perf_counter() - th0) * 1e3 / args.steps     # host time to QUEUE a step (no sync inside)
    e1.record(st)
    barrier()
    seg_report("resident loop", seg[-args.steps:])
    if dbg_losses and rank == 0:
        print("[losses] warm-up + resident loop:", [round(float(v), 3) for v in dbg_losses], file=sys.stderr)
    del seg[:], dbg_losses[:]
    launches = int(lib.tt_kernel_launches() - k0)
    if gstep is not None:
        launches = int(gstep.launches_per_replay(*resident[0]) or 0) * args.steps      # replays do not pass through the host counter
    elif towers is not None:
        launches += int(towers.launches_per_replay(*resident[0]) or 0) * args.steps     # eager part counted above + the two replays
    sampler.mark_end()
    clocks = sampler.stop()
    ms_total = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    ms_step = ms_total.item() / args.steps
    value = world * B / (ms_step * 1e-3)
    # ---- e2e: pinned host ids -> device every step, loss read back every step, wall clock, max over ranks
    # (the next step's ids are copied on a copy stream while the current step computes - still one H2D per step
    #  inside the timed region)
    copy_stream = torch.cuda.Stream(dev)
    # double-buffered device id slots, allocated once (a fresh allocation on the copy stream every step made the caching
    # allocator's cross-stream bookkeeping part of the timed region: occasional 100+ ms outliers)
    slots = [(torch.empty_like(resident_shape[0], device=dev), torch.empty_like(resident_shape[1], device=dev)) for _ in range(2)]
    slot_free = [None, None]                 # event: the step that read the slot has been queued and has finished

    def upload(i):
        q, d = host[i % nb]
        qd, dd = slots[i & 1]
        with torch.cuda.stream(copy_stream):
            if slot_free[i & 1] is not None:
                copy_stream.wait_event(slot_free[i & 1])
            qd.copy_(q, non_blocking=True)
            dd.copy_(d, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return qd, dd, ev

    # the loss of step i goes to a pinned host slot with an async copy right after the step is queued and is READ on
    # the host while step i+1 runs (one-step-late logging, as a training loop would do it): one D2H read per step
    # inside the timed region, but the host never drains the launch queue between steps
    LAG = max(1, int(os.environ.get("TT_E2E_LAG", "1")))          # how many steps late a loss is read (measurement knob)
    NS = LAG + 1
    loss_host = torch.empty(NS, dtype=torch.float32).pin_memory()
    # blocking events: the host thread sleeps in synchronize() instead of spinning (8 ranks share the box's cores with their
    # NCCL / sampler threads)
    loss_ev = [torch.cuda.Event(blocking=True) for _ in range(NS)]
    losses = []
    dbg = os.environ.get("TT_E2E_DEBUG") == "1"
    marks, step_done = [], []
    sampler2 = ClockSampler(local)
    sampler2.start()
    reset_training_state()
    for i in range(args.warmup):
        step(*resident[i % nb])
    torch.cuda.synchronize()
    time.sleep(0.2)
    ee0, ee1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler2.mark_begin()
    t0 = time.perf_counter()
    ee0.record(st)
    nxt = upload(0)
    t_up0 = time.perf_counter()
    for i in range(args.steps):
        ta = time.perf_counter()
        qd, dd, ev = nxt
        if i + 1 < args.steps:
            nxt = upload(i + 1)
        st.wait_event(ev)
        tb = time.perf_counter()
        loss, _ = step(qd, dd)
        slot_free[i & 1] = torch.cuda.Event()
        slot_free[i & 1].record(st)
        tc = time.perf_counter()
        if dbg:
            step_done.append(torch.cuda.Event(enable_timing=True))
            step_done[-1].record(st)
        loss_host[i % NS].copy_(loss.detach(), non_blocking=True)
        loss_ev[i % NS].record(st)
        if i >= LAG:
            loss_ev[(i - LAG) % NS].synchronize()
            losses.append(float(loss_host[(i - LAG) % NS]))
        if dbg:
            marks.append((round((tb - ta) * 1e3, 2), round((tc - tb) * 1e3, 2), round((time.perf_counter() - tc) * 1e3, 2)))
    t_loop = time.perf_counter()
    for j in range(max(0, args.steps - LAG), args.steps):
        loss_ev[j % NS].synchronize()
        losses.append(float(loss_host[j % NS]))
    if dbg and rank == 0:
        print("e2e ms: first upload %.2f, loop %.2f, last loss %.2f" % ((t_up0 - t0) * 1e3, (t_loop - t_up0) * 1e3,
                                                                        (time.perf_counter() - t_loop) * 1e3), file=sys.stderr)
        print("e2e per-step host ms (upload, launch of the step, loss copy + wait for the previous loss):", marks, file=sys.stderr)
        seg_report("e2e loop", seg[-args.steps:])
        print("[losses] warm-up + e2e loop:", [round(float(v), 3) for v in dbg_losses], file=sys.stderr)
        torch.cuda.synchronize()
        tk = [r[3].elapsed_time(r[4]) for r in seg[-args.steps:]]
        print("[mining top-k ms per e2e step]", [round(v, 2) for v in tk], file=sys.stderr)
        torch.cuda.synchronize()
        print("e2e device ms between consecutive step ends:", [round(a.elapsed_time(b), 2) for a, b in zip(step_done, step_done[1:])],
              file=sys.stderr)
    loss_val = losses[-1]
    ee1.record(st)
    t_b = time.perf_counter()
    barrier()
    e2e_wall = time.perf_counter() - t0
    sampler2.mark_end()
    clocks_e2e = sampler2.stop()
    e2e_dev_ms = ee0.elapsed_time(ee1) / args.steps
    if dbg and rank == 0:
        print("e2e ms: closing barrier %.2f" % ((time.perf_counter() - t_b) * 1e3), file=sys.stderr)
    e2e_s = torch.tensor([e2e_wall], device=dev)
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_val = world * B * args.steps / e2e_s.item()
    # ---- dominant-kernel timing (LAST: its eager passes change allocator / library state, and the two timed loops above
    # must see the same state): the same K steps again with the towers on ONE stream (kernel durations under
    # two-stream overlap would include time-slicing with the other tower), CUDA events around every launch
    model.overlap_towers = False
    lib.tt_debug_wgrad_overlap(0)
    eager_step(*resident[0])
    barrier()
    lib.tt_prof_enable(1)
    for i in range(args.steps):
        eager_step(*resident[i % nb])
    barrier()
    roofs, fams = kernel_rooflines(lib, args.steps, B, peaks())
    gemm_rows, gemm_roof = gemm_table(lib, args.steps, peaks())
    lib.tt_prof_enable(0)
    lib.tt_debug_wgrad_overlap(0 if args.serial_towers else 1)
    model.overlap_towers = not args.serial_towers
    h2d = B * (TQ + TD) * 8
    if rank == 0:
        pk = peaks()
        line = {
            "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 accumulate; forward GEMMs as 3 fp16 hi/lo tensor passes, backward GEMMs 1 fp16 pass",
            "data": "synthetic", "config": workload_config(B, world, V),
            "e2e": {"value": e2e_val, "unit": "pairs/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "device_ms_per_step": e2e_dev_ms, "clocks": clocks_e2e,
                    "note": "wall clock over K steps through the nn.Module API: pinned int64 ids -> device on a copy stream "
                            "(one step ahead), every step's loss copied to pinned memory and read on the host one step late"},
            "gpu_launches": launches, "host_queue_ms_per_step": host_launch_ms, "clocks": clocks, "loss": loss_val,
            "engine": lib.tt_engine().decode(),
            "step_tflops_per_gpu": FLOPS_PER_PAIR * B / (ms_step * 1e-3) / 1e12,
            "step_frac_of_bf16_sustained": FLOPS_PER_PAIR * B / (ms_step * 1e-3) / 1e12 / pk["bf16_sustained"],
        }
        line["roofline"] = roofs[0] if roofs else None
        line["kernels"] = roofs[1:] + ([gemm_roof] if gemm_roof else [f for f in fams[2:]])
        line["gemm_table"] = gemm_rows
        line["graph"] = {"captured": ("whole step" if gstep is not None else "towers (forward + backward); loss, NCCL, optimizer eager"
                                      if towers is not None else False),
                         "launches_per_replay": (gstep.launches_per_replay(*resident[0]) if gstep is not None else
                                                 towers.launches_per_replay(*resident[0]) if towers is not None else None),
                         "note": "host_queue_ms_per_step is the host time to queue one step: input copies + one cudaGraphLaunch when "
                                 "captured, every kernel launch through Python/ctypes otherwise (--no-graph)"}
    # ---- parity self-check on this very process group (small shapes, ~1 s): the sharded step - eager and captured in a CUDA
    # graph, NCCL included - against one process on the concatenated batch: loss, every gradient, mined indices
    from two_towers_b200.selfcheck import dp_parity_check
    try:
        pc = dp_parity_check(group, dev, graphed=graphed)
        parity = {"status": "ok" if pc["ok"] else "MISMATCH", "world": world, "loss_rel": pc["loss_rel"],
                  "worst_grad_rel": pc["worst_grad"][1], "worst_grad": pc["worst_grad"][0], "indices_equal": pc["indices_equal"],
                  "graphed_weight_rel": pc.get("graphed_weight_rel"), "failed": pc.get("failed"),
                  "what": "data-parallel step vs single process on the concatenated batch (two_towers_b200/selfcheck.py)"}
    except Exception as exc:                                  # a failing check must not hide the measurement; it is reported
        parity = {"status": "ERROR", "error": repr(exc)[:300]}
    if rank == 0:
        line["parity_check"] = parity
    del resident[:], host[:]
    del model, opt, gstep, towers, towers_step, step, eager_step, reducer
    torch.cuda.empty_cache()
    if rank == 0 and world == 1 and not args.no_configs:
        line["configs"] = secondary_configs(args, dev, table, graphed)
    del table
    torch.cuda.empty_cache()
    ev = eval_leg(args, dev, rank, world, group) if args.eval_queries > 0 else None
    if rank == 0:
        line["eval"] = ev
        if world == 1 and not args.no_torch_gpu:
            try:
                rows = torch_gpu_arm(args, 5, 3)
                line["torch_gpu_baseline"] = [{"value": r["value"], "unit": "pairs/s", "ms_per_step": r["ms_per_step"], "dtype": r["dtype"],
                                               "ratio_ours_over_it": value / r["value"]} for r in rows]
            except Exception as exc:
                line["torch_gpu_baseline"] = {"error": repr(exc)[:300]}
        if world == 1 and not args.no_cpu_baseline:
            val, ms, threads = cpu_reference(3, 1, args.cpu_batch)
            line["cpu_baseline"] = {"value": val, "unit": "pairs/s", "cores": threads, "kind": "port",
                                    "sample": f"{args.cpu_batch} pairs/step, same model/loss/mining/optimizer, "
                                              f"pre-embedded fp32 inputs, 3 timed steps after 1 warm-up ({ms:.0f} ms/step)"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def workload_config(B, world, V):
    return {"workload": "enhanced_two_tower: EnhancedTwoTowerModel(300,128) bi-GRU h256 x2 layers + LayerNorm head, "
                        "MarginRankingLoss in-batch (global negatives under DP), row top-8 hard negatives, Adam step "
                        "(BASELINE configs[2] at 1 GPU, configs[3] at N GPUs)",
            "batch_per_gpu": B, "global_batch": B * world, "query_len": TQ, "doc_len": TD, "vocab": V,
            "parallelism": f"dp{world}", "l2": "activations per step (>10 GB) exceed the 126 MB L2; 4 rotating id batches"}


if __name__ == "__main__":
    a = parse()
    {"reference": run_reference, "torch-gpu": run_torch_gpu, "ours": run_ours}[a.impl](a)
