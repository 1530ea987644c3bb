"""GPU: the allocation-free entry point tt_opt_step_noalloc (scratch in a file-scope `static __device__` array; no
counterpart in the reference) against the existing tt_opt_step with a
caller workspace, through the C ABI, on two calls in a row (the second with FEWER chunks than the first, so stale scratch
contents from the first call are in the array), plus the size limit.  Collected last on purpose (file name).
Passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

from two_towers_b200 import _lib as L
from two_towers_b200 import optim as O

pytestmark = pytest.mark.gpu
MAX_CHUNKS = 16384                  # TT_OPT_NOALLOC_MAX_CHUNKS (include/tt_b200.h)


def _state(dev, shapes, seed):
    g = torch.Generator().manual_seed(seed)
    mk = lambda s, sc: (torch.randn(*s, generator=g) * sc).to(dev)
    return [dict(p=mk(s, 0.3), g=mk(s, 2.0), m=mk(s, 0.01), v=mk(s, 0.01).abs()) for s in shapes]


def _clone(st):
    return [{k: t.clone() for k, t in e.items()} for e in st]


def _args(st, dev, norm, workspace, step):
    tb, cb, nchunks = O.build_tables([(e["p"].data_ptr(), e["g"].data_ptr(), e["m"].data_ptr(), e["v"].data_ptr(), e["p"].numel())
                                      for e in st])
    tens, chunks = tb.to(dev), cb.to(dev)
    b1, b2 = 0.9, 0.999
    a = O.OptStep(tens.data_ptr(), chunks.data_ptr(), len(st), nchunks,
                  O.OPT_NORM | O.OPT_STEP | O.OPT_CLIP | O.OPT_SKIP_NONFINITE, 1,
                  1e-3, b1, b2, 1e-8, 0.01, 1.0, 1.0 - b1 ** step, (1.0 - b2 ** step) ** 0.5,
                  norm.data_ptr(), workspace.data_ptr() if workspace is not None else None,
                  workspace.numel() if workspace is not None else 0, None, None)
    return a, (tens, chunks), nchunks


def test_opt_step_noalloc_matches_opt_step_on_two_calls_in_a_row():
    dev = torch.device("cuda", 0)
    lib = L.lib()
    stream = torch.cuda.current_stream(dev).cuda_stream
    # call 1: 11 chunks (two tensors span 3, odd sizes, a scalar); call 2: 4 chunks - the scratch still holds call 1's sums
    for call, shapes in enumerate([[(768, 256), (70001,), (3,), (1,), (256, 513), (65536,)], [(300, 128), (65537,), (5,)]]):
        ref = _state(dev, shapes, 10 + call)
        got = _clone(ref)
        n_ref = torch.zeros(4, device=dev)
        n_got = torch.zeros(4, device=dev)
        a_ref, keep_ref, nchunks = _args(ref, dev, n_ref, None, call + 1)
        ws = L.workspace(lib.tt_opt_step_workspace(nchunks), dev)
        a_ref.workspace, a_ref.workspace_bytes = ws.data_ptr(), ws.numel()
        L.check(lib.tt_opt_step(C.byref(a_ref), stream), "tt_opt_step")
        a_got, keep_got, _ = _args(got, dev, n_got, None, call + 1)          # no workspace at all
        L.check(lib.tt_opt_step_noalloc(C.byref(a_got), stream), "tt_opt_step_noalloc")
        torch.cuda.synchronize(dev)
        assert nchunks == (11 if call == 0 else 4)
        assert n_ref[0].item() > 1.0 and n_ref[1].item() < 1.0               # the clip is active: the norm matters
        # same kernels, same order of summation: the tolerance of the optimizer tests (2e-6 relative) is met with equality
        assert torch.allclose(n_got, n_ref, rtol=2e-6, atol=0.0), (n_got, n_ref)
        for e_got, e_ref in zip(got, ref):
            for k in ("p", "g", "m", "v"):
                assert torch.allclose(e_got[k], e_ref[k], rtol=2e-6, atol=2e-7), (call, k)


def test_opt_step_noalloc_rejects_sizes_beyond_the_static_scratch():
    dev = torch.device("cuda", 0)
    lib = L.lib()
    st = _state(dev, [(8,)], 1)
    norm = torch.zeros(4, device=dev)
    a, keep, _ = _args(st, dev, norm, None, 1)
    before = st[0]["p"].clone()
    a.nchunks = MAX_CHUNKS + 1
    rc = lib.tt_opt_step_noalloc(C.byref(a), torch.cuda.current_stream(dev).cuda_stream)
    torch.cuda.synchronize(dev)
    assert rc != 0
    msg = lib.tt_last_error().decode()
    assert "TT_OPT_NOALLOC_MAX_CHUNKS" in msg and str(MAX_CHUNKS) in msg, msg
    assert torch.equal(st[0]["p"], before)                                    # nothing was launched
