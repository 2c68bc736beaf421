#!/usr/bin/env python3
"""brute-force top-2 (BASELINE config 5, 2048 x 1 048 576) pairs/s on one GPU, device-resident; checks the result checksum"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "orb-slam2_b200")]
import numpy as np, torch, orbb200, synth
dev = torch.device("cuda", 0)
L = orbb200.lib()
NQ, NDB = 2048, 1 << 20
q, db, _ = synth.descriptor_db(NQ, NDB, seed=7)
qd, dbd = torch.from_numpy(q).to(dev), torch.from_numpy(db).to(dev)
out = torch.empty((NQ, 4), dtype=torch.int32, device=dev)
ts = torch.cuda.Stream(device=dev); torch.cuda.set_stream(ts); st = ts.cuda_stream
def step(): orbb200.check(L.orbb200_match_bruteforce_top2_device(qd.data_ptr(), NQ, dbd.data_ptr(), NDB, out.data_ptr(), 0, st))
for _ in range(3): step()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record()
for _ in range(10): step()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
r = out.cpu().numpy()
print(json.dumps({"lib": os.environ.get("ORBB200_LIB_PATH", "default"), "ms": ms, "pairs_per_s": NQ * NDB / (ms * 1e-3), "checksum": int(r[:, 1].astype(np.int64).sum()), "dsum": int(r[:, 0].sum() + r[:, 2].sum())}))
