"""One table for the Snake matcher parity tests: per matcher, the seeded case generator (snake_cases.py), a digest of
its inputs, and three runners — the reference's own compiled code (`ref`, oracle/_ref/libsnake_ref.so), the CPU
restatement (`oracle`) and the CUDA library through the C ABI (`gpu`). Every runner returns the same list of arrays."""
import numpy as np

import snake_cases as SC

I32 = np.int32


def _fine_out(r):
    m, k2p, bi, bd, vis, pts = r[:6]
    bd = np.where(bi >= 0, bd, -1)  # the distance of an unmatched point is unspecified
    return [np.array([m], I32), k2p, bi, bd.astype(I32), vis, pts["valid"].astype(I32)]


def _rect(O, c):
    return O.undistort_keypoints(c["rl"], c["left"])[0], O.undistort_keypoints(c["rr"], c["right"])[0]


def _v01(c):
    return (c["v1"] == 1).astype(np.uint8), (c["v2"] == 1).astype(np.uint8)


def _loop_fuse_state(c, bi, bd):
    from test_oracle_snake_live import fuse_loop_edits

    fused, state, bad = fuse_loop_edits(c, bi, bd)
    return [np.array([fused], I32), state, bad]


def _bow_loop(c, one):
    """MatchBoW candidate by candidate -> the batch's output form (n_matches[n_cand], match12[n_cand, n1])."""
    nm, rows = [], []
    for d2, fv2, v2 in c["cands"]:
        m, m12 = one(d2, fv2, (v2 == 1).astype(np.uint8))
        nm.append(m), rows.append(np.asarray(m12, I32)[:len(c["d1"])])
    return [np.array(nm, I32), np.stack(rows).astype(I32)]


F32_12 = float(np.float32(1.2))

MATCHERS = {
    # name: (case generator, input arrays for the digest, ref runner, oracle runner, gpu runner)
    "stereo": (
        SC.case_stereo,
        lambda c: (c["left"], c["dl"], c["right"], c["dr"]),
        lambda O, c: list(O.ref_stereo_matching(c["rl"], c["rr"], c["bf"], c["relaxed"], SC.LEVELS, F32_12, c["left"], c["dl"],
                                                c["right"], c["dr"])),
        lambda O, c: list(O.stereo_match(O.stereo_params(c["bf"], c["relaxed"], SC.SCALES), *sum(
            [[k, d] for k, d in zip(_rect(O, c), (c["dl"], c["dr"]))], []))[:3]),
        lambda so, ctx, c: list(ctx.StereoMatchingRaw(c["rl"], c["rr"], so.stereo_params(c["bf"], c["relaxed"], SC.SCALES),
                                                      c["left"], c["dl"], c["right"], c["dr"])),
    ),
    "rgbd": (
        SC.case_rgbd,
        lambda c: (c["kp"], c["depth"]),
        lambda O, c: list(O.ref_stereo_from_rgbd(c["p"], c["kp"], c["depth"])),
        lambda O, c: list(O.stereo_from_rgbd(c["p"], c["kp"], c["depth"])),
        lambda so, ctx, c: list(ctx.ComputeStereoFromRGBD(c["p"], c["kp"], c["depth"])),
    ),
    "fine": (
        SC.case_fine,
        lambda c: (c["f"].kp, c["f"].desc, c["pts"]),
        lambda O, c: _fine_out(O.ref_search_projection_fine(c["f"].view, c["pts"], c["cam"], SC.pyramid(), c["th"], c["ratio"])),
        lambda O, c: _fine_out(O.search_projection_fine(c["f"].view, c["pts"], c["cam"], SC.pyramid(), c["th"], c["ratio"])),
        lambda so, ctx, c: _fine_out(ctx.SearchByProjection2(c["f"].view, c["pts"], c["cam"], SC.pyramid(), c["th"], c["ratio"])),
    ),
    "coarse": (
        SC.case_coarse,
        lambda c: (c["f"].kp, c["f"].desc, c["pts"]),
        lambda O, c: list(O.ref_search_projection_coarse(c["f"].view, c["pts"], c["cam"], SC.pyramid(), c["th"], c["fe"], c["motion"])),
        lambda O, c: list(O.search_projection_coarse(c["f"].view, c["pts"], c["cam"], SC.pyramid(), c["th"], c["fe"], c["motion"])[:2]),
        lambda so, ctx, c: list(ctx.SearchByProjectionFrameFrame2(c["f"].view, c["pts"], c["cam"], SC.pyramid(), c["th"], c["fe"],
                                                                  c["motion"])),
    ),
    "frame_to_kf": (
        SC.case_kf,
        lambda c: (c["f"].kp, c["f"].desc, c["pts"]),
        lambda O, c: list(O.ref_search_projection_frame_to_kf(c["f"].view, c["pts"], c["cam"], c["th"], c["fe"])),
        lambda O, c: list(O.search_projection_frame_to_kf(c["f"].view, c["pts"], c["cam"], c["th"], c["fe"])[:2]),
        lambda so, ctx, c: list(ctx.SearchByProjectionFrameToKeyframe(c["f"].view, c["pts"], c["cam"], c["th"], c["fe"])),
    ),
    "match_bow": (
        SC.case_bow,
        lambda c: (c["d1"], c["d2"], c["v1"], c["v2"]) + c["fv1"]._keep + c["fv2"]._keep,
        lambda O, c: list(O.ref_match_bow(c["d1"], c["d2"], c["fv1"], c["fv2"], c["v1"], c["v2"], c["th"], c["ratio"])),
        lambda O, c: list(O.match_bow(c["d1"], c["d2"], c["fv1"], c["fv2"], *_v01(c), c["th"], c["ratio"])[:2]),
        lambda so, ctx, c: list(ctx.MatchBoW(c["d1"], c["d2"], c["fv1"], c["fv2"], *_v01(c), c["th"], c["ratio"])),
    ),
    "match_bow_batch": (
        SC.case_bow_batch,
        lambda c: (c["d1"], c["v1"]) + c["fv1"]._keep + tuple(a for d2, fv2, v2 in c["cands"] for a in (d2, v2) + fv2._keep),
        lambda O, c: list(O.ref_match_bow_batch(c["d1"], c["fv1"], c["v1"], c["cands"], c["th"], c["ratio"])),
        lambda O, c: _bow_loop(c, lambda d2, fv2, v2: O.match_bow(c["d1"], d2, c["fv1"], fv2, (c["v1"] == 1).astype(np.uint8), v2,
                                                                  c["th"], c["ratio"])[:2]),
        lambda so, ctx, c: list(ctx.MatchBoWBatch(c["d1"], c["fv1"], (c["v1"] == 1).astype(np.uint8),
                                                  [(d2, fv2, (v2 == 1).astype(np.uint8)) for d2, fv2, v2 in c["cands"]], c["th"],
                                                  c["ratio"])),
    ),
    "mapping_fuse": (
        SC.case_fuse,
        lambda c: (c["f"].kp, c["f"].desc, c["pts"]),
        lambda O, c: list(O.ref_fuse_search(c["f"].view, c["mask"], c["pts"], c["cam"], SC.pyramid(), c["th"], c["obs"], c["fth"])),
        lambda O, c: list(O.fuse_search(c["f"].view, c["mask"], c["pts"], c["cam"], SC.pyramid(), c["th"], c["obs"], c["fth"])[:2]),
        lambda so, ctx, c: list(ctx.Fuse(c["f"].view, c["mask"], c["pts"], c["cam"], SC.pyramid(), c["th"], c["obs"], c["fth"])),
    ),
    "loop_fuse": (
        SC.case_fuse_loop,
        lambda c: (c["f"].kp, c["f"].desc, c["pts"], c["skip"]),
        lambda O, c: list(O.ref_fuse_loop(c["f"].view, c["skip"], c["pts"], c["cam"], min(c["th"], 4.0))),
        lambda O, c: _loop_fuse_state(c, *O.fuse_search_loop(c["f"].view, c["skip"], c["pts"], c["cam"], min(c["th"], 4.0))),
        lambda so, ctx, c: _loop_fuse_state(c, *ctx.FuseLoopSearch(c["f"].view, c["skip"], c["pts"], c["cam"], min(c["th"], 4.0))),
    ),
    "triangulation_bow": (
        SC.case_tri_bow,
        lambda c: (c["f1"].kp, c["f1"].desc, c["f2"].kp, c["f2"].desc, c["E"]),
        lambda O, c: [O.ref_search_triangulation_bow(c["f1"].view, c["np1"], c["fv1"], c["f2"].view, c["np2"], c["fv2"], c["E"],
                                                     c["cam"], c["epi"], c["fd"])],
        lambda O, c: [O.search_triangulation_bow(c["f1"].view, c["np1"], c["fv1"], c["f2"].view, c["np2"], c["fv2"], c["E"],
                                                 c["cam"], c["epi"], c["fd"])[0]],
        lambda so, ctx, c: [ctx.SearchForTriangulation2(c["f1"].view, c["np1"], c["fv1"], c["f2"].view, c["np2"], c["fv2"], c["E"],
                                                        c["cam"], c["epi"], c["fd"])],
    ),
    "triangulation_project": (
        SC.case_tri_project,
        lambda c: (c["f1"].kp, c["f1"].desc, c["f2"].kp, c["f2"].desc, c["E"], c["grid"]),
        lambda O, c: [O.ref_search_triangulation_project(c["f1"].view, c["np1"], c["f2"].view, c["np2"], c["grid"], c["E"],
                                                         c["cam"], c["epi"], c["fd"])],
        lambda O, c: [O.search_triangulation_project(c["f1"].view, c["np1"], c["f2"].view, c["np2"], c["grid"], c["E"],
                                                     c["cam"], c["epi"], c["fd"])[0]],
        lambda so, ctx, c: [ctx.SearchForTriangulationProject(c["f1"].view, c["np1"], c["f2"].view, c["np2"], c["grid"], c["E"],
                                                              c["cam"], c["epi"], c["fd"])],
    ),
    "prediction": (
        SC.case_prediction,
        lambda c: (c["f"].kp, c["f"].desc, c["dl"], c["pos"]),
        lambda O, c: list(O.ref_match_prediction(c["dl"], c["pos"], c["f"].view, 50.0, 50)),
        lambda O, c: list(O.match_prediction(c["dl"], c["pos"], c["f"].view, 50.0, 50)),
        lambda so, ctx, c: list(ctx.ComputeMatchesPrediction(c["dl"], c["pos"], c["f"].view, 50.0, 50)),
    ),
}

# the frozen cases: (seed, big) — small tie-heavy ones and a few at BASELINE sizes (C2/C3/C4-like)
GOLDEN_CASES = [(s, False) for s in range(40)] + [(s, True) for s in (1000, 1001, 1002)]


def norm(outs):
    """Runner output -> list of contiguous arrays (scalars become 1-element int32 arrays)."""
    res = []
    for o in outs:
        a = np.asarray(o)
        if a.ndim == 0:
            a = a.reshape(1).astype(I32)
        res.append(np.ascontiguousarray(a))
    return res


def same(a, b):
    return len(a) == len(b) and all(x.shape == y.shape and x.tobytes() == y.tobytes() for x, y in zip(a, b))
