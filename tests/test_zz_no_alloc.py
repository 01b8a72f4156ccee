"""dba_damping_gather_noalloc (include/dba.h): the damping gather with its frame -> row map in the library's static
device array instead of a cudaMalloc'ed buffer.  No counterpart in the reference;
the check is against the existing entry point dba_damping_gather, which tests/test_glue_gpu.py pins to
the restatement of factor_graph.pyc update@227.  Collected last on purpose (file name).
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, N, ht, wd, E, drop):
    g = torch.Generator().manual_seed(seed)
    ii = torch.randint(0, N, (E,), generator=g)
    for f in drop:                                            # frames that are not a source: rows must skip them
        ii = ii[ii != f]
    damping = torch.rand((N, ht, wd), generator=g) * 1e-3
    return damping.cuda(), ii.cuda()


def test_noalloc_matches_damping_gather_twice_in_a_row():
    from cut3r_slam_b200 import glue
    # two calls in a row with DIFFERENT graphs and frame counts: the second must not see the first call's map
    cases = [_case(11, 40, 48, 64, 300, drop=(3, 7, 39)), _case(12, 23, 43, 77, 90, drop=(0, 11))]
    refs = [glue.damping_eta(d, ii) for d, ii in cases]       # existing entry point (allocates through the context)
    torch.cuda.synchronize()
    outs = [torch.full_like(r, float("nan")) for r in refs]   # the caller's outputs, allocated before the calls
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for (d, ii), out in zip(cases, outs):
        glue.damping_eta_noalloc(d, ii, out)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before            # nothing allocated through torch on this path
    for out, ref in zip(outs, refs):
        assert out.shape == ref.shape
        assert torch.equal(out, ref)                          # same kernels, same arithmetic: bit-identical
    # and the first graph again after the second (the shared array keeps the last call's contents)
    again = torch.full_like(refs[0], float("nan"))
    glue.damping_eta_noalloc(cases[0][0], cases[0][1], again)
    torch.cuda.synchronize()
    assert torch.equal(again, refs[0])


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    from cut3r_slam_b200 import droid_backends as db
    ctx = db.context(torch.device("cuda:0"))
    d = torch.zeros(8, device="cuda")                         # never dereferenced: the size check comes first
    ii = torch.zeros(4, dtype=torch.int64, device="cuda")
    out = torch.zeros(8, device="cuda")
    rc = ctx.lib.dba_damping_gather_noalloc(ctx.handle, C.c_void_p(d.data_ptr()), C.c_void_p(ii.data_ptr()), 4, 12001,
                                            1, 1, 1, C.c_void_p(out.data_ptr()), C.c_void_p(0))
    assert rc < 0
    assert b"12000" in ctx.lib.dba_last_error(ctx.handle)
    # the limit itself is accepted (N = 12000 frames of 1x1 pixels)
    N = 12000
    damping = torch.rand((N, 1, 1), generator=torch.Generator().manual_seed(5)).cuda()
    src = torch.arange(0, N, 3, device="cuda")
    from cut3r_slam_b200 import glue
    ref = glue.damping_eta(damping, src)
    out = torch.empty_like(ref)
    glue.damping_eta_noalloc(damping, src, out)
    torch.cuda.synchronize()
    assert torch.equal(out, ref)
