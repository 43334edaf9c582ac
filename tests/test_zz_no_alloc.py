"""The scratch-free vote-scoring entry point, exercised through the C ABI.

``fm_vote_scores_static`` exists for hosts that cannot give the library scratch memory because they
cannot allocate on the device when they need a score (no ``cudaMalloc`` after start-up, a GPU whose
memory another framework already owns, a bare C host).  It must return exactly what
``fm_vote_scores`` returns with caller-provided scratch -- on the first call, and again on a second
call that reuses the built-in array with different, smaller inputs -- and must refuse sizes beyond
its stated limits.  This file is named so that pytest collects it last.
"""
import ctypes as C

import pytest
import torch

from fedmse_decentralized_b200.layout import DEFAULT_LAYOUT as L
from fedmse_decentralized_b200.ops import _lib

pytestmark = pytest.mark.gpu
TILE = 64


def _tiles(n_sel, rows, dev):
    t = [(p, s, min(TILE, rows - s), p * rows + s) for p in range(n_sel) for s in range(0, rows, TILE)]
    return torch.tensor(t, dtype=torch.int32, device=dev)


def _case(seed, n_local, n_sel, rows, dev):
    g = torch.Generator().manual_seed(seed)
    slab = torch.zeros(n_local, L.padded)
    slab[:, :L.n_params] = (torch.rand(n_local, L.n_params, generator=g) - 0.5) * 0.3
    V = torch.zeros(rows, 128)
    V[:, :115] = torch.randn(rows, 115, generator=g) * 3 + 1
    sel = torch.randperm(n_local, generator=g)[:n_sel].to(torch.int32)
    jit = 1.0 + (torch.rand(n_sel, generator=g) - 0.5) * 2e-4
    return slab.to(dev), V.to(dev), sel.to(dev), jit.to(dev), _tiles(n_sel, rows, dev)


def _with_caller_scratch(lib, stream, slab, V, sel, jit, tiles, voter, chunk):
    n_sel, rows, dev = sel.numel(), V.shape[0], V.device
    norm = torch.empty(rows, 128, device=dev)
    sse = torch.empty(n_sel * rows, device=dev)
    out = torch.full((n_sel,), -1.0, device=dev)
    a = _lib.FmVoteArgs(slab.data_ptr(), sel.data_ptr(), n_sel, 0, slab.shape[0], V.data_ptr(), rows, chunk,
                        jit.data_ptr(), voter, tiles.data_ptr(), tiles.shape[0], norm.data_ptr(), sse.data_ptr(),
                        out.data_ptr())
    _lib.check(lib.fm_vote_scores(C.byref(a), stream), "fm_vote_scores")
    return out


def _with_builtin_scratch(lib, stream, slab, V, sel, jit, tiles, voter, chunk):
    n_sel, rows = sel.numel(), V.shape[0]
    out = torch.full((n_sel,), -1.0, device=V.device)
    a = _lib.FmVoteStaticArgs(slab.data_ptr(), sel.data_ptr(), n_sel, 0, slab.shape[0], V.data_ptr(), rows, chunk,
                              jit.data_ptr(), voter, tiles.data_ptr(), tiles.shape[0], out.data_ptr())
    rc = lib.fm_vote_scores_static(C.byref(a), stream)
    return rc, out


def test_builtin_scratch_matches_caller_scratch_twice_in_a_row():
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    # call 1 fills the built-in array almost to its limits; call 2 is smaller, so stale contents of
    # call 1 sit right behind (and inside the footprint of) what call 2 uses
    for seed, n_local, n_sel, rows, voter in ((1, 70, _lib.STATIC_VOTE_MAX_SEL, 1000, 3), (2, 9, 5, 167, 0)):
        slab, V, sel, jit, tiles = _case(seed, n_local, n_sel, rows, dev)
        want = _with_caller_scratch(lib, stream, slab, V, sel, jit, tiles, voter, 128)
        rc, got = _with_builtin_scratch(lib, stream, slab, V, sel, jit, tiles, voter, 128)
        assert rc == 0, lib.fm_vote_scores_static_error().decode()
        torch.cuda.synchronize()
        assert float(got[voter]) == 0.0 and bool((got[torch.arange(n_sel) != voter] > 0).all())
        assert torch.allclose(got, want, rtol=1e-4, atol=1e-6), (got - want).abs().max()
        assert torch.equal(got, want)     # same kernels, same order: bit-identical


def test_builtin_scratch_rejects_oversize_requests():
    dev = torch.device("cuda", 0)
    lib = _lib.load()
    stream = torch.cuda.current_stream().cuda_stream
    slab, V, sel, jit, tiles = _case(3, 4, 3, _lib.STATIC_VOTE_MAX_ROWS + 1, dev)
    rc, _ = _with_builtin_scratch(lib, stream, slab, V, sel, jit, tiles, 0, 128)
    assert rc != 0
    msg = lib.fm_vote_scores_static_error().decode()
    assert "exceeds the built-in scratch" in msg and str(_lib.STATIC_VOTE_MAX_ROWS) in msg
    slab, V, sel, jit, tiles = _case(4, 80, _lib.STATIC_VOTE_MAX_SEL + 1, 64, dev)
    rc, _ = _with_builtin_scratch(lib, stream, slab, V, sel, jit, tiles, 0, 128)
    assert rc != 0 and "models" in lib.fm_vote_scores_static_error().decode()
    torch.cuda.synchronize()
