#!/usr/bin/env python3
"""Single-frame / single-pair latency of the drop-in (BASELINE config 1 as written: ONE 640x480 frame, TUM mono settings; and the
KITTI stereo pair on two host threads, src/Frame.cc:82-85), through the C++ facade (orb-slam2_b200/facade_extract.bin --latency),
with the CUDA graph on and off, beside the reference's own ORBextractor / stereo Frame constructor on one / two host threads
(oracle/_ref/liborbslam_ref.so).  Prints one JSON object; bench.py embeds it as `latency`."""
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle")]
import synth  # noqa: E402


def facade(w, h, nfeat, left, right, iters, graph):
    exe = os.path.join(ROOT, "orb-slam2_b200", "facade_extract.bin")
    r = subprocess.run([exe, "--latency", str(w), str(h), str(nfeat), "8", left, right, str(iters), str(graph)], capture_output=True, text=True,
                       timeout=600)
    if r.returncode != 0:
        return {"error": (r.stderr or r.stdout)[-300:]}
    return json.loads(r.stdout.strip().splitlines()[-1])


def cpu_reference(w, h, nfeat, imgs, iters):
    import orbslam_ref as R
    if not R.available() or iters <= 0:
        return None
    R.set_monotonic(0)  # plain malloc: timing, not tie-break pinning
    us = []
    if len(imgs) == 1:
        ex = R.Extractor(nfeat, 1.2, 8, 20, 7)
        for it in range(iters + 2):
            t0 = time.perf_counter()
            ex.extract(imgs[0])
            if it >= 2:
                us.append((time.perf_counter() - t0) * 1e6)
        what = "reference ORBextractor::operator(), one host thread (image calls = plain-C primitives, not OpenCV SIMD)"
    else:
        for it in range(iters + 2):
            t0 = time.perf_counter()
            f = R.RefFrame(imgs[0], imgs[1], nfeat, 1.2, 8, 20, 7)
            if it >= 2:
                us.append((time.perf_counter() - t0) * 1e6)
            del f
        what = "reference stereo Frame constructor: ExtractORB on two host threads + ComputeStereoMatches + grid (Frame.cc:65-125)"
    R.set_monotonic(1)
    us.sort()
    return {"call": what, "p50_us": us[len(us) // 2], "p99_us": us[min(len(us) - 1, int(len(us) * 0.99))], "iters": iters}


def run(iters=300, cpu_iters=15):
    out = {}
    with tempfile.TemporaryDirectory() as td:
        tum = synth.textured_frame(640, 480, 1000)
        L, Rr = synth.stereo_pair(1241, 376, 1001)
        p = {}
        for name, a in (("tum", tum), ("kl", L), ("kr", Rr)):
            p[name] = os.path.join(td, name + ".raw")
            a.tofile(p[name])
        out["tum_640x480_1000feat_mono"] = {"graph": facade(640, 480, 1000, p["tum"], "-", iters, 1), "pipeline": facade(640, 480, 1000, p["tum"], "-", iters, 0),
                                            "cpu_reference": cpu_reference(640, 480, 1000, [tum], cpu_iters)}
        out["kitti_1241x376_2000feat_stereo_pair"] = {"graph": facade(1241, 376, 2000, p["kl"], p["kr"], iters, 1),
                                                       "pipeline": facade(1241, 376, 2000, p["kl"], p["kr"], iters, 0),
                                                       "cpu_reference": cpu_reference(1241, 376, 2000, [L, Rr], cpu_iters)}
    return out


if __name__ == "__main__":
    print(json.dumps(run()))
