"""CPU: the C-ABI library loads and exports every symbol include/*.h declares (tt_b200.h: the boundary; tt_b200_debug.h:
test / measurement hooks); struct layouts seen by
ctypes equal the C compiler's.  No compute calls (no GPU here)."""
import ctypes
import os
import re
import subprocess
import tempfile

import pytest

from conftest import ROOT
from two_towers_b200 import _lib

HEADER = os.path.join(ROOT, "include", "tt_b200.h")
HEADERS = [HEADER, os.path.join(ROOT, "include", "tt_b200_debug.h")]


def _declared_symbols(headers=HEADERS):
    out = set()
    for h in headers:
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        out |= set(re.findall(r"\b(tt_[a-z0-9_]+)\s*\(", src))
    return sorted(out)


def test_debug_hooks_are_not_in_the_boundary_header():
    public = _declared_symbols([HEADER])
    assert not [s for s in public if s.startswith("tt_debug_") or s.startswith("tt_gemm_")], public


def test_header_declares_functions():
    syms = _declared_symbols()
    assert len(syms) >= 20 and "tt_gru_layer_fwd" in syms and "tt_rows_topk" in syms


def test_library_exports_every_declared_symbol():
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for s in _declared_symbols():
        assert hasattr(handle, s), f"{s} declared in include/*.h but not exported"


def test_binding_covers_every_declared_symbol():
    assert sorted(_lib.SIGNATURES) == _declared_symbols()


def test_library_reports_version_and_engine_without_gpu():
    lib = _lib.lib()
    assert lib.tt_abi_version() == 2
    assert lib.tt_engine().decode() in ("simt-fp32", "tcgen05-bf16x3", "tcgen05-f16x3-fwd/f16-bwd")


def test_struct_layouts_match_c_compiler():
    prog = r'''
#include <stdio.h>
#include <stddef.h>
#include "tt_b200.h"
int main(void){
  printf("%zu %zu %zu %zu\n", sizeof(tt_gru_layer), sizeof(tt_gru_layer_grad), sizeof(tt_head), sizeof(tt_head_grad));
  printf("%zu %zu %zu %zu %zu\n", offsetof(tt_gru_layer,x), offsetof(tt_gru_layer,seed), offsetof(tt_gru_layer,w_ih),
         offsetof(tt_gru_layer,lengths), offsetof(tt_gru_layer,workspace_bytes));
  printf("%zu %zu %zu\n", offsetof(tt_head,seed), offsetof(tt_head,v_lo), offsetof(tt_head,norm));
  printf("%zu %zu\n", offsetof(tt_gru_layer_grad,d_out), offsetof(tt_head_grad,d_out));
  printf("%zu %zu %zu %zu %zu %zu\n", sizeof(tt_gru_stack), sizeof(tt_gru_stack_grad), offsetof(tt_gru_stack,x),
         offsetof(tt_gru_stack,seed), offsetof(tt_gru_stack,saved_bytes), offsetof(tt_gru_stack_grad,d_weights));
  printf("%zu %zu %zu %zu\n", offsetof(tt_gru_stack,seed_dev), offsetof(tt_head,seed_dev), offsetof(tt_opt_args,step_dev), offsetof(tt_opt_args,consts_dev));
  printf("%zu %zu %zu %zu %zu %zu %zu\n", sizeof(tt_opt_tensor), sizeof(tt_opt_chunk), sizeof(tt_opt_args), offsetof(tt_opt_chunk,offset),
         offsetof(tt_opt_args,lr), offsetof(tt_opt_args,bias_correction2_sqrt), offsetof(tt_opt_args,workspace_bytes));
  return 0; }
'''
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "l.c")
        open(c, "w").write(prog)
        exe = os.path.join(d, "l")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        out = subprocess.check_output([exe]).decode().split()
    out = list(map(int, out))
    G, GG, H, HG = _lib.GruLayer, _lib.GruLayerGrad, _lib.Head, _lib.HeadGrad
    exp = [ctypes.sizeof(G), ctypes.sizeof(GG), ctypes.sizeof(H), ctypes.sizeof(HG),
           G.x.offset, G.seed.offset, G.w_ih.offset, G.lengths.offset, G.workspace_bytes.offset,
           H.seed.offset, H.v_lo.offset, H.norm.offset, GG.d_out.offset, HG.d_out.offset]
    S, SG = _lib.GruStack, _lib.GruStackGrad
    exp += [ctypes.sizeof(S), ctypes.sizeof(SG), S.x.offset, S.seed.offset, S.saved_bytes.offset, SG.d_weights.offset]
    from two_towers_b200 import optim as O
    exp += [S.seed_dev.offset, H.seed_dev.offset, O.OptStep.step_dev.offset, O.OptStep.consts_dev.offset]
    exp += [ctypes.sizeof(O.OptTensor), ctypes.sizeof(O.OptChunk), ctypes.sizeof(O.OptStep), O.OptChunk.offset.offset,
            O.OptStep.lr.offset, O.OptStep.bias_correction2_sqrt.offset, O.OptStep.workspace_bytes.offset]
    assert out == exp


def test_cpu_tensors_are_rejected_loudly():
    import torch
    import two_towers_b200 as tt
    m = tt.EnhancedTwoTowerModel(20, 4)
    with pytest.raises(RuntimeError, match="CUDA"):
        m(torch.randn(2, 3, 20), torch.randn(2, 3, 20))
    with pytest.raises(RuntimeError, match="CUDA"):
        tt.InfoNCELoss()(torch.randn(2, 4), torch.randn(2, 4))
    with pytest.raises(RuntimeError, match="CUDA"):
        tt.get_hard_negatives(torch.randn(4), torch.randn(5, 4), 0, 2)


def test_retrieval_size_queries_are_host_only_and_consistent():
    """tt_corpus_index_bytes / tt_rows_topk*_workspace / tt_debug_topk_fallback_flags are pure host arithmetic: callable
    without a GPU, monotonic in the problem size, and the indexed scan (corpus part in the caller's index buffer) never
    needs more scratch than the scan that converts the corpus on every call."""
    lib = _lib.lib()
    D = 128
    sizes = [lib.tt_corpus_index_bytes(n, D) for n in (0, 1000, 100_000, 1_100_000)]
    assert sizes == sorted(sizes) and sizes[0] >= 256
    assert sizes[-1] >= 1_100_000 * 4                      # at least the row norms
    for mode in (0, 1):
        for q, n in ((1, 1000), (300, 20_000), (8192, 1_100_000)):
            plain = lib.tt_rows_topk_workspace(q, n, D, mode)
            indexed = lib.tt_rows_topk_indexed_workspace(q, n, D, mode)
            assert 0 < indexed <= plain
            for k in (1, 10):
                for flag in (0, 1):
                    off = lib.tt_debug_topk_fallback_flags(q, n, D, mode, k, flag)
                    limit = indexed if flag else plain
                    assert off == -1 or (0 <= off and off + 4 * q <= limit)
    with pytest.raises(Exception):
        _lib.check(lib.tt_debug_sel_passes(7), "tt_debug_sel_passes")
    assert lib.tt_debug_sel_passes(2) == 0 and lib.tt_debug_sel_resident(1) == 0 and lib.tt_debug_sel_prefetch(4) == 0
