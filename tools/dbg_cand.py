import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb-slam2_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orbb200, ref_cpu, synth
w, h = 640, 480
img = synth.textured_frame(w, h, 1000)
ref = ref_cpu.RefExtractor(1000, 1.2, 8, 20, 7)
ref.extract(img)
ex = orbb200.ORBextractor(1000, 1.2, 8, 20, 7)
ex(img)
for l in range(8):
    c = ref.level_candidates(l); g = ex.debug_candidates(l)
    n = min(len(c), len(g))
    bad = np.nonzero((g[:n, 0] != c["x"][:n]) | (g[:n, 1] != c["y"][:n]) | (g[:n, 2] != c["resp"][:n]))[0]
    print("level", l, "n", len(g), len(c), "mismatches", len(bad))
    for i in bad[:12]:
        print("   i", i, "gpu", g[i], "ref", c["x"][i], c["y"][i], c["resp"][i], " x%31=", (g[i][0] - 3) % 31, "y%30", (g[i][1] - 3) % 30)
