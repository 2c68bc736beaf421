#!/usr/bin/env python3
"""Per-call time, pairs/s and captures/s of the frame-sized searches (host buffers in, host results out -- the call a tracker makes):
window search (SearchByProjection shape: 2000 windows on a 2000-feature frame), candidate distances / top-2 (SearchByBoW shape),
SearchForTriangulation (2000 x ~50), stereo matching of one KITTI pair.  One JSON object on stdout."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2_b200")]
import numpy as np, torch, orbb200, synth

def timed(fn, reps=200, warm=10):
    for _ in range(warm): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps

res = {}
rng = np.random.default_rng(3)
m = orbb200.ORBmatcher()
# --- windowed search (KITTI frame, 2000 features, 2000 windows of radius 15 * scale) -------------------------------------------------
w, h = 1241, 376
img = synth.textured_frame(w, h, 11)
ex = orbb200.ORBextractor(2000, 1.2, 8, 20, 7)
kps, desc = ex(img)
nq = 2000
src = rng.integers(0, len(kps), nq)
win = np.zeros((nq, 5), np.float32)
win[:, 0] = kps["x"][src] + rng.normal(0, 3, nq); win[:, 1] = kps["y"][src] + rng.normal(0, 3, nq)
win[:, 2] = 15.0 * np.float32(1.2) ** kps["octave"][src]
win[:, 3] = kps["octave"][src] - 1; win[:, 4] = kps["octave"][src] + 1
grid = (0.0, 0.0, 64.0 / w, 48.0 / h, 64, 48)
qd = desc[src].copy()
offs, idx, dist = m.window_search(kps, desc, grid, win, qd, cap=64 * nq)
t = timed(lambda: m.window_search(kps, desc, grid, win, qd, cap=64 * nq))
res["window_search"] = {"queries": nq, "targets": int(len(kps)), "pairs": int(offs[-1]), "us_per_call": t * 1e6, "pairs_per_s": offs[-1] / t,
                        "calls_per_s": 1 / t}
# --- candidate lists (SearchByBoW shape: 2000 queries x ~12 candidates) -----------------------------------------------------------------
q, tdb, _ = synth.descriptor_db(2000, 2000, seed=5)
sizes = rng.poisson(12, 2000)
co = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
ci = rng.integers(0, 2000, int(co[-1])).astype(np.int32)
t = timed(lambda: m.candidate_distances(q, tdb, co, ci))
res["candidate_distances"] = {"queries": 2000, "pairs": int(co[-1]), "us_per_call": t * 1e6, "pairs_per_s": co[-1] / t, "calls_per_s": 1 / t}
t = timed(lambda: m.match_candidates(q, tdb, co, ci))
res["match_candidates"] = {"queries": 2000, "pairs": int(co[-1]), "us_per_call": t * 1e6, "pairs_per_s": co[-1] / t, "calls_per_s": 1 / t}
# --- SearchForTriangulation (2000 x ~50) -------------------------------------------------------------------------------------------------
nt = 2000
tkp = np.zeros(nt, orbb200.KP_DTYPE)
tkp["x"], tkp["y"] = rng.uniform(0, w, nt).astype(np.float32), rng.uniform(0, h, nt).astype(np.float32)
tkp["octave"] = rng.integers(0, 8, nt)
sizes = rng.poisson(50, 2000)
so = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
si = rng.integers(0, nt, int(so[-1])).astype(np.int32)
qxy = np.stack([rng.uniform(0, w, 2000), rng.uniform(0, h, 2000)], 1).astype(np.float32)
st0 = np.zeros(2000, np.uint8)
F12 = np.array([[0, 0, 0], [0, 0, -1], [0, 1, 0]], np.float32)
sf = np.float32(1.2) ** np.arange(8, dtype=np.float32); sig2 = (sf * sf).astype(np.float32)
t = timed(lambda: m.search_triangulation(q, qxy, st0, tdb, tkp, st0, so, si, F12, 600.0, 180.0, sf, sig2))
res["search_triangulation"] = {"queries": 2000, "pairs": int(so[-1]), "us_per_call": t * 1e6, "pairs_per_s": so[-1] / t, "calls_per_s": 1 / t}
# --- ComputeStereoMatches of one KITTI pair (host keypoints + device pyramids of the two extractors) ------------------------------------
L_img, R_img = synth.stereo_pair(w, h, 5)[:2]
exL, exR = orbb200.ORBextractor(2000, 1.2, 8, 20, 7), orbb200.ORBextractor(2000, 1.2, 8, 20, 7)
exL(L_img); exR(R_img)
try:
    t = timed(lambda: orbb200.stereo_match(exL, exR, 386.1448, 0.5372), reps=100)
    res["stereo_match_pair"] = {"us_per_call": t * 1e6, "calls_per_s": 1 / t}
except Exception as e:  # the probe reports, the tests check
    res["stereo_match_pair"] = {"error": str(e)[:200]}
print(json.dumps(res))
