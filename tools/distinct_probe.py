"""Measures MapPoint::ComputeDistinctiveDescriptors batched over map points (orbb200_distinctive_descriptors): the kernel
alone with the groups resident in HBM (CUDA events on the launch stream), the host-buffer call (H2D + kernel + D2H) and the
oracle port on one host core (how LocalMapping runs the reference loop).  Prints one JSON line."""
import ctypes
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402
import orbb200  # noqa: E402
import ref_cpu  # noqa: E402
import distinct_cases as dc  # noqa: E402


def main():
    npoints = int(os.environ.get("DD_POINTS", "200000"))
    reps = int(os.environ.get("DD_REPS", "20"))
    rng = np.random.default_rng(5)
    # observation counts of a mature map: most points seen by a handful of keyframes, a tail seen by dozens
    sizes = np.minimum(2 + rng.geometric(0.12, npoints), 120).astype(np.int64)
    offs = np.zeros(npoints + 1, np.int32)
    offs[1:] = np.cumsum(sizes)
    total = int(offs[-1])
    base = rng.integers(0, 256, (npoints, 32), dtype=np.uint8)
    desc = np.repeat(base, sizes, axis=0)
    noise = (rng.random((total, 256)) < 0.06)
    desc ^= np.packbits(noise, axis=1)
    pairs = int((sizes.astype(np.int64) ** 2).sum())
    m = orbb200.ORBmatcher()
    L = m.L
    d_desc = torch.from_numpy(desc).cuda()
    d_offs = torch.from_numpy(offs).cuda()
    d_idx = torch.zeros(npoints, dtype=torch.int32, device="cuda")
    d_med = torch.zeros(npoints, dtype=torch.int32, device="cuda")
    ts = torch.cuda.Stream()
    torch.cuda.set_stream(ts)
    vp = ctypes.c_void_p
    flush = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")
    ms = []
    for it in range(reps + 3):
        flush.add_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        orbb200.check(L.orbb200_distinctive_descriptors_device(vp(d_desc.data_ptr()), vp(d_offs.data_ptr()), npoints,
                                                               vp(d_idx.data_ptr()), vp(d_med.data_ptr()), vp(ts.cuda_stream)))
        e1.record()
        torch.cuda.synchronize()
        if it >= 3:
            ms.append(e0.elapsed_time(e1))
    k_ms = float(np.median(ms))
    gi, gm = m.distinctive_descriptors(desc, offs)
    t0 = time.perf_counter()
    for _ in range(3):
        gi, gm = m.distinctive_descriptors(desc, offs)
    host_ms = (time.perf_counter() - t0) / 3 * 1e3
    ns = min(npoints, 20000)
    t0 = time.perf_counter()
    ri, rm = ref_cpu.distinctive_descriptors(desc[:offs[ns]], offs[:ns + 1])
    cpu_s = time.perf_counter() - t0
    out = {"probe": "MapPoint::ComputeDistinctiveDescriptors batched", "points": npoints, "descriptors": total,
           "mean_observations": float(sizes.mean()), "max_observations": int(sizes.max()), "distance_pairs_NxN": pairs,
           "kernel_ms_L2_flushed": k_ms, "kernel_points_per_s": npoints / (k_ms * 1e-3), "kernel_pairs_per_s": pairs / (k_ms * 1e-3),
           "algorithmic_bytes": total * 32 + (npoints + 1) * 4 + npoints * 8,
           "kernel_GBps_algorithmic": (total * 32 + (npoints + 1) * 4 + npoints * 8) / (k_ms * 1e-3) / 1e9,
           "host_call_ms": host_ms, "host_call_points_per_s": npoints / (host_ms * 1e-3),
           "cpu_port_points_per_s_1_thread": ns / cpu_s, "cpu_sample_points": ns,
           "parity_device": bool(np.array_equal(d_idx.cpu().numpy()[:ns], ri) and np.array_equal(d_med.cpu().numpy()[:ns], rm)),
           "parity_host_call": bool(np.array_equal(gi[:ns], ri) and np.array_equal(gm[:ns], rm))}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
