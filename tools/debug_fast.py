import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import orbb200, ref_cpu, synth
img = synth.textured_frame(640, 480, 1000)
ref = ref_cpu.RefExtractor(1000, 1.2, 8, 20, 7)
ref.extract(img)
ex = orbb200.ORBextractor(1000, 1.2, 8, 20, 7)
ex(img)
RING = [(0,3),(1,3),(2,2),(3,1),(3,0),(3,-1),(2,-2),(1,-3),(0,-3),(-1,-3),(-2,-2),(-3,-1),(-3,0),(-3,1),(-2,2),(-1,3)]
def score(im, x, y):
    v = int(im[y, x]); d = [v - int(im[y+dy, x+dx]) for dx, dy in RING]; d = d + d
    best = 0
    for k in range(16):
        a = d[k:k+9]; best = max(best, min(a), -max(a))
    return best - 1
c = ref.level_candidates(0); g = ex.debug_candidates(0)
print("cpu", c[:8]); print("gpu", g[:8])
for (x, y) in [(3,5),(3,7),(7,7),(5,9)]:
    print((x,y), "numpy score", score(img, x+16, y+16))
    print(img[y+16-3:y+16+4, x+16-3:x+16+4])
