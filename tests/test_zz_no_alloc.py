"""mslam_b200_match_knn_ratio_noalloc_dev (include/mslam_b200.h): the matcher for callers that cannot allocate -- no context,
the kernels' scratch in a static __device__ array of the library.  No counterpart in the reference.
Called through the C ABI and checked, bit for bit (indices and Hamming distances are integers: the
tolerance of every matcher test here is zero), against the existing entry point mslam_b200_match_knn_ratio, on calls in a row
that share the scratch array.  This file sorts last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _sets(rng, n_from, n_to, planted):
    """`to` = random descriptors, `planted` of them noisy copies of `from` rows (so that the ratio test keeps some)"""
    frm = rng.integers(0, 256, (n_from, 32), dtype=np.uint8)
    to = rng.integers(0, 256, (n_to, 32), dtype=np.uint8)
    for q in rng.choice(n_to, size=min(planted, n_to), replace=False):
        bits = np.unpackbits(frm[rng.integers(0, n_from)])
        bits ^= (rng.random(256) < 0.04).astype(np.uint8)
        to[q] = np.packbits(bits)
    return frm, to


def _noalloc(pkg, torch, frm, to, ratio=0.7):
    n_from, n_to = len(frm), len(to)
    d_from = torch.from_numpy(np.ascontiguousarray(frm)).cuda() if n_from else torch.empty(32, dtype=torch.uint8, device="cuda")
    d_to = torch.from_numpy(np.ascontiguousarray(to)).cuda() if n_to else torch.empty(32, dtype=torch.uint8, device="cuda")
    m = max(n_to, 1)
    o_f = torch.full((m,), -7, dtype=torch.int32, device="cuda")
    o_t = torch.full((m,), -7, dtype=torch.int32, device="cuda")
    o_d = torch.full((m,), -7, dtype=torch.int32, device="cuda")
    o_n = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    st = torch.cuda.current_stream().cuda_stream
    pkg.match_knn_ratio_noalloc_dev(d_from, n_from, d_to, n_to, o_f, o_t, o_d, o_n, ratio=ratio, device=torch.cuda.current_device(),
                                    stream=st)
    torch.cuda.synchronize()
    n = int(o_n.item())
    return o_f[:n].cpu().numpy(), o_t[:n].cpu().numpy(), o_d[:n].cpu().numpy()


def test_noalloc_matcher_equals_the_context_matcher_on_calls_in_a_row(pkg):
    import torch
    rng = np.random.default_rng(20261105)
    # a large pair first, then smaller and ragged ones: stale keys of an earlier call must never be read; sizes cover the
    # sliced (train set >= 256) and unsliced forms, the 64-query CTA seam, and the maximum
    cases = [(1000, 1000, 600), (1000, 1000, 600), (300, 77, 40), (65, 1000, 300), (pkg.NOALLOC_MAX_DESC, pkg.NOALLOC_MAX_DESC, 900),
             (2, 5, 2), (513, 129, 64)]
    with pkg.Context(max_batch=1, nfeatures=500, max_descriptors=pkg.NOALLOC_MAX_DESC) as ctx:
        for i, (nf, nt, planted) in enumerate(cases):
            frm, to = _sets(rng, nf, nt, planted)
            if i == 1:  # the same sizes again with other data
                frm = frm[::-1].copy()
            fi, ti, di = ctx.match_knn_ratio(frm, to)
            gf, gt, gd = _noalloc(pkg, torch, frm, to)
            assert len(gf) == len(fi), f"case {i}: {len(gf)} matches vs {len(fi)}"
            assert planted < 10 or len(fi) >= planted // 2, f"case {i}: the generator gave only {len(fi)} matches"
            assert (gf == fi.astype(np.int64)).all() and (gt == ti.astype(np.int64)).all() and (gd == di).all(), f"case {i}"
        # the other ratio the tests use
        frm, to = _sets(rng, 700, 700, 300)
        with pkg.Context(max_batch=1, nfeatures=500, max_descriptors=1024, ratio=0.8) as c8:
            fi, ti, di = c8.match_knn_ratio(frm, to)
        gf, gt, gd = _noalloc(pkg, torch, frm, to, ratio=0.8)
        assert (gf == fi.astype(np.int64)).all() and (gt == ti.astype(np.int64)).all() and (gd == di).all()


def test_noalloc_matcher_edge_cases_and_limits(pkg):
    import torch
    rng = np.random.default_rng(7)
    frm, to = _sets(rng, 40, 30, 10)
    # the reference is undefined for an empty `to` or fewer than two `from` descriptors: contract = no match, count written
    for f, t in ((frm[:1], to), (frm[:0], to), (frm, to[:0])):
        gf, gt, gd = _noalloc(pkg, torch, f, t)
        assert len(gf) == 0
    # beyond the static scratch: rejected with a message, nothing launched
    big = torch.zeros((pkg.NOALLOC_MAX_DESC + 1, 32), dtype=torch.uint8, device="cuda")
    out = torch.zeros(pkg.NOALLOC_MAX_DESC + 1, dtype=torch.int32, device="cuda")
    n = torch.zeros(1, dtype=torch.int32, device="cuda")
    for nf, nt in ((pkg.NOALLOC_MAX_DESC + 1, 10), (10, pkg.NOALLOC_MAX_DESC + 1)):
        with pytest.raises(pkg.MslamB200Error) as ei:
            pkg.match_knn_ratio_noalloc_dev(big, nf, big, nt, out, out, out, n)
        assert ei.value.status == pkg.E_INVALID and "MSLAM_B200_NOALLOC_MAX_DESC" in str(ei.value)
    with pytest.raises(pkg.MslamB200Error):
        pkg.match_knn_ratio_noalloc_dev(big, 10, big, 10, out, out, out, None)
    # and it still works afterwards
    fi_ref = None
    with pkg.Context(max_batch=1, nfeatures=500, max_descriptors=64) as ctx:
        fi_ref = ctx.match_knn_ratio(frm, to)
    gf, gt, gd = _noalloc(pkg, torch, frm, to)
    assert (gf == fi_ref[0].astype(np.int64)).all() and (gt == fi_ref[1].astype(np.int64)).all() and (gd == fi_ref[2]).all()
