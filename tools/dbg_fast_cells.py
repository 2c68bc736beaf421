"""debug aid: which grid cells of a level differ between the CUDA FAST stage and the oracle"""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle")]
import numpy as np, orbb200, ref_cpu, synth
w, h, nf = 1241, 376, 2000
img = synth.textured_frame(w, h, 1000)
ref = ref_cpu.RefExtractor(nf, 1.2, 8, 20, 7); ref.extract(img)
ex = orbb200.ORBextractor(nf, 1.2, 8, 20, 7); ex(img)
for l in range(8):
    c = ref.level_candidates(l); g = ex.debug_candidates(l)
    sc = set((int(a["x"]), int(a["y"]), int(a["resp"])) for a in c); sg = set(map(tuple, g.tolist()))
    miss = sorted(sc - sg, key=lambda t: (t[1], t[0])); extra = sorted(sg - sc, key=lambda t: (t[1], t[0]))
    print("level", l, "oracle", len(c), "gpu", len(g), "missing", len(miss), "extra", len(extra))
    if miss: print("  missing:", miss[:12])
    if extra: print("  extra:", extra[:12])
