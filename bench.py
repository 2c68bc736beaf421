#!/usr/bin/env python
"""bench.py -- ORB front-end hot path on B200: KITTI-shape stereo extraction (+ ComputeStereoMatches) frames/s,
and brute-force Hamming top-2 match pairs/s, with HBM / POPC rooflines and the CPU port timed beside them.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--pairs P] [--impl b200|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
         bench.py --gpus N --steps K --warmup W

One step = one pass of the hot path over one batch of P synthetic KITTI-shape stereo pairs per GPU (2P images of
1241x376, 2000 features, 8 levels, FAST 20/7): pyramid -> FAST cells -> quadtree -> blur -> orientation +
descriptors for all 2P images, then Frame::ComputeStereoMatches for the P pairs.  Frames are sharded over the ranks
(weak scaling, no data-path collective).  `value` = images/s with inputs resident in HBM; `e2e` = images/s through
the host-buffer C-ABI (orbb200_extract_stereo_batch) with pinned host inputs/outputs and
all H2D/D2H copies inside the timed region.  The `match` object is BASELINE config 5 (2 048 queries x 1 048 576
database rows, database row-sharded over the ranks, NCCL all-gather of the per-shard top-2 + merge kernel).

PyTorch is plumbing only here (device buffers, pinned memory, torch.distributed); every kernel launched inside
the timed regions is one of liborbb200.so's.  oracle/ (the CPU port) is executed only for `cpu_baseline` and for
`--impl reference`.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

KITTI = dict(w=1241, h=376, nfeatures=2000, scale=1.2, nlevels=8, ini=20, min=7, bf=386.1448, fx=718.856)
METRIC = "ORB extract frames/s @KITTI 1241x376 2k feats"
WORKLOAD = "kitti_stereo_1241x376_L+R_2000feat_8lvl_extract+ComputeStereoMatches"
NQ, NDB = 2048, 1 << 20


def level_sizes(w, h, scale, nlevels):
    """pyramid level shapes exactly as ORBextractor.cc:450,459,1351-1353 (f32 running product, cvRound)."""
    sf = [np.float32(1.0)]
    for _ in range(1, nlevels):
        sf.append(np.float32(np.float64(sf[-1]) * np.float64(np.float32(scale))))
    out = []
    for s in sf:
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
    return out


def algorithmic_bytes(w, h, scale, nlevels, nkp):
    """per-stage algorithmic bytes of one extracted image."""
    P = [a * b for a, b in level_sizes(w, h, scale, nlevels)]
    tot = sum(P)
    stages = {
        "pyramid": sum(P[l - 1] + P[l] for l in range(1, nlevels)),
        "fast_cells": tot,
        "octree_select": 0,
        "blur7": 2 * tot,
        "orient_describe": nkp * (749 + 512 + 32 + 28),
    }
    return stages, sum(stages.values())


def make_pairs(npairs, seed0=1000, nbase=6):
    """npairs stereo pairs [2*npairs, H, W] u8, interleaved L,R.  A few seeded base pairs (synth.stereo_pair) are
    expanded by row-rolls, which keep the per-row disparity structure and give distinct pixel data per pair."""
    import synth
    w, h = KITTI["w"], KITTI["h"]
    base = [synth.stereo_pair(w, h, seed0 + k) for k in range(min(nbase, npairs))]
    out = np.empty((2 * npairs, h, w), np.uint8)
    for p in range(npairs):
        L, R = base[p % len(base)]
        sh = (p // len(base)) * 13
        out[2 * p] = np.roll(L, sh, axis=0)
        out[2 * p + 1] = np.roll(R, sh, axis=0)
    return out


class ClockSampler(threading.Thread):
    """samples SM clock + throttle reasons of one GPU through NVML while the timed region runs"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# --------------------------------------------------------------------------------------------- CPU arms
def make_config(P):
    """the workload both arms are quoted on (identical keys and values in the b200 and the reference line)"""
    w, h = KITTI["w"], KITTI["h"]
    B = 2 * P
    return {"workload": WORKLOAD, "pairs_per_step_per_gpu": P, "images_per_step_per_gpu": B, "input_bytes_per_step_per_gpu": B * w * h,
            "l2_policy": "inputs larger than L2 (126 MB)" if B * w * h > 126e6 else "inputs smaller than L2: raise --pairs",
            "sharding": "frames per GPU, no collective"}


def cpu_reference_pairs(imgs, nthreads):
    """the REFERENCE's own code (oracle/_ref/liborbslam_ref.so = unmodified ORBextractor.cc / Frame.cc compiled against the OpenCV
    stand-in): per pair the stereo Frame constructor -- ExtractORB of both images on two threads, ComputeStereoMatches, grid --
    with persistent extractors, one session per host thread.  Returns (seconds, images)."""
    import orbslam_ref as R
    from concurrent.futures import ThreadPoolExecutor
    K = KITTI
    R.set_monotonic(0)  # plain malloc (the monotonic arena only matters for tie-break pinning in the tests)
    npairs = len(imgs) // 2
    tl = threading.local()

    def work(p):
        if not hasattr(tl, "s"):
            tl.s = R.Session(K["nfeatures"], K["scale"], K["nlevels"], K["ini"], K["min"], bf=K["bf"])
        return tl.s.frame(imgs[2 * p], imgs[2 * p + 1])[0]

    with ThreadPoolExecutor(nthreads) as ex:
        list(ex.map(work, range(min(nthreads, npairs))))  # warm: build the per-thread sessions
        t0 = time.perf_counter()
        list(ex.map(work, range(npairs)))
        dt = time.perf_counter() - t0
    R.set_monotonic(1)
    return dt, 2 * npairs


def cpu_extract_pairs(imgs, nthreads):
    """the plain-C restatement (oracle/orb_ref.c) on host threads: extraction of every image + ComputeStereoMatches per
    pair; returns (seconds, images).  ctypes releases the GIL inside the C calls."""
    import ref_cpu
    from concurrent.futures import ThreadPoolExecutor
    K = KITTI
    npairs = len(imgs) // 2
    b = K["bf"] / K["fx"]
    tl = threading.local()

    def work(p):
        if not hasattr(tl, "ex"):
            tl.ex = (ref_cpu.RefExtractor(K["nfeatures"], K["scale"], K["nlevels"], K["ini"], K["min"]),
                     ref_cpu.RefExtractor(K["nfeatures"], K["scale"], K["nlevels"], K["ini"], K["min"]))
        eL, eR = tl.ex
        kL, dL = eL.extract(imgs[2 * p])
        kR, dR = eR.extract(imgs[2 * p + 1])
        pyrL = [eL.level_image(l) for l in range(K["nlevels"])]
        pyrR = [eR.level_image(l) for l in range(K["nlevels"])]
        tb = eL.tables()
        ref_cpu.stereo_match(kL, dL, kR, dR, pyrL, pyrR, tb["scale"], tb["inv_scale"], K["bf"], b)
        return len(kL)

    with ThreadPoolExecutor(nthreads) as ex:
        list(ex.map(work, range(min(nthreads, npairs))))  # warm: build per-thread extractors
        t0 = time.perf_counter()
        list(ex.map(work, range(npairs)))
        dt = time.perf_counter() - t0
    return dt, 2 * npairs


def cpu_library_view(img):
    """the OpenCV calls the reference delegates to (src/ORBextractor.cc:988, 1314, 1366), timed with the OpenCV on
    this image (cv2, SIMD) on ONE host thread for one KITTI image: 7 resizes, 8 blurs, FAST per 30-px cell on every level."""
    try:
        import cv2
    except Exception:
        return None
    cv2.setNumThreads(1)
    sizes = level_sizes(KITTI["w"], KITTI["h"], KITTI["scale"], KITTI["nlevels"])
    det = cv2.FastFeatureDetector_create(KITTI["ini"], True)
    out = {}
    reps = 3
    t0 = time.perf_counter()
    for _ in range(reps):
        pyr = [img]
        for (w, h) in sizes[1:]:
            pyr.append(cv2.resize(pyr[-1], (w, h), interpolation=cv2.INTER_LINEAR))
    out["resize_ms"] = (time.perf_counter() - t0) / reps * 1e3
    t0 = time.perf_counter()
    for _ in range(reps):
        for m in pyr:
            cv2.GaussianBlur(m, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    out["blur_ms"] = (time.perf_counter() - t0) / reps * 1e3
    t0 = time.perf_counter()
    for _ in range(reps):
        for m in pyr:
            h, w = m.shape
            for y0 in range(16, h - 16 - 6, 30):
                for x0 in range(16, w - 16 - 6, 30):
                    det.detect(m[y0:min(y0 + 36, h - 16), x0:min(x0 + 36, w - 16)])
    out["fast_cells_ms"] = (time.perf_counter() - t0) / reps * 1e3
    out["note"] = "cv2 %s, one thread, one 1241x376 image; the quadtree, orientation, descriptors and stereo matching are the reference's own loops and are not part of this view" % cv2.__version__
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on the box's host cores (oracle/_ref, compiled from
    the unmodified sources; OpenCV calls served by the stand-in over plain-C primitives), same config / metric as the b200 arm,
    each step a bounded sample of the step's pairs."""
    if rank != 0:
        return
    import orbslam_ref as R
    cores = os.cpu_count() or 1
    P = args.pairs
    have_ref = R.available()
    nthreads = max(1, cores // 2) if have_ref else cores  # the stereo constructor runs two extraction threads per session
    npairs = min(P, max(nthreads, 4) * 8)                # bounded sample of the step's P pairs (about 0.5 s wall per step)
    imgs = make_pairs(P)[:2 * npairs]
    fn = cpu_reference_pairs if have_ref else cpu_extract_pairs
    for _ in range(max(1, min(args.warmup, 2))):
        fn(imgs[:2 * min(npairs, nthreads)], nthreads)
    tot_t, tot_img = 0.0, 0
    for _ in range(args.steps):
        dt, n = fn(imgs, nthreads)
        tot_t += dt
        tot_img += n
    v = tot_img / tot_t
    kind = "reference" if have_ref else "port"
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": make_config(P),
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": kind,
                         "sample": "%d of the step's %d stereo pairs per step x %d steps; %d sessions x 2 extraction threads; "
                                   "%s" % (npairs, P, args.steps, nthreads,
                                           "unmodified src/ORBextractor.cc + Frame.cc stereo constructor, OpenCV calls through the plain-C stand-in"
                                           if have_ref else "oracle/orb_ref.c port")},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------- B200 arm
def run_b200(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import orbb200
    import synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    K = KITTI
    P = args.pairs
    B = 2 * P
    w, h = K["w"], K["h"]
    bconst = K["bf"] / K["fx"]
    # host side of the feed: this rank's threads (and the pinned buffers it first-touches below) on the CPUs local to its GPU
    bound_cpus = 0 if args.no_bind else orbb200.lib().orbb200_bind_host_to_device(local_rank)
    ex = orbb200.ORBextractor(K["nfeatures"], K["scale"], K["nlevels"], K["ini"], K["min"], device=local_rank)
    cap = ex.cap
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)

    # ---- inputs: pinned host copy (e2e) and device-resident copy (value); different pairs on every rank
    host_imgs = torch.empty((B, h, w), dtype=torch.uint8).pin_memory()
    host_imgs.numpy()[:] = make_pairs(P, seed0=1000 + 97 * rank)
    dev_imgs = host_imgs.to(dev)
    kps_d = torch.empty((B, cap, 7), dtype=torch.float32, device=dev)
    desc_d = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    n_d = torch.zeros((B,), dtype=torch.int32, device=dev)
    uR_d = torch.empty((P, cap), dtype=torch.float32, device=dev)
    dep_d = torch.empty((P, cap), dtype=torch.float32, device=dev)
    nv_d = torch.zeros((P,), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    def step_device():
        ex.extract_batch_device(dev_imgs.data_ptr(), B, w, h, w, w * h, kps_d.data_ptr(), desc_d.data_ptr(), cap, n_d.data_ptr())
        ex.stereo_match_batch_device(P, K["bf"], bconst, uR_d.data_ptr(), dep_d.data_ptr(), nv_d.data_ptr())

    for _ in range(max(args.warmup, 3)):
        step_device()
    ex.sync()
    launches_per_step = ex.last_launches() + 2
    ex.set_profiling(True)  # stage events on the handle's stream, inside the timed region
    sampler = ClockSampler(local_rank)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.start()
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    ex.sync()
    barrier()
    sampler.stop_flag = True
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    stage_sum, ncalls = ex.stage_ms_sum()
    ex.set_profiling(False)
    ms_per_step = ms_total / args.steps
    value = world * B * args.steps / (ms_total * 1e-3)
    n_host = n_d.cpu().numpy()
    nv_host = nv_d.cpu().numpy()
    mean_kp = float(n_host.mean())

    # ---- roofline of the dominant kernel (per-stage CUDA-event time inside the timed region)
    stage_ms = {k: v / max(ncalls, 1) for k, v in stage_sum.items()}
    # all stages run back to back on the handle's stream: stereo = step - whole extraction (first launch to last)
    extract_ms = ex.extract_total_ms_sum / max(ncalls, 1)
    stage_ms["stereo_match"] = max(ms_per_step - extract_ms, 0.0) if ncalls == args.steps else None
    stage_ms["extract_first_to_last_launch"] = extract_ms
    abytes, abytes_img = algorithmic_bytes(w, h, K["scale"], K["nlevels"], mean_kp)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    dom = max((k for k in ex.STAGES), key=lambda k: stage_ms[k])
    dom_launches = (K["nlevels"] - 1) if dom == "pyramid" else 1
    dom_bytes = abytes[dom] * B  # algorithmic bytes all launches of that stage move in one step
    dom_bytes_note = None
    if dom == "octree_select":  # no image traffic: candidates in (4 B) + selected out (4 B)
        dom_bytes = 0
        dom_bytes_note = "quadtree selection moves no image bytes; it is latency/serial bound (see DESIGN.md)"
    achieved = dom_bytes / (stage_ms[dom] * 1e-3) / 1e9 if stage_ms[dom] > 0 else 0.0
    # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel, from this round's committed `ncu --set full`
    # capture of the same command (profiles/r2_traffic.json, written by tools/ncu_traffic.py); null when there is no capture
    # for this kernel and batch -- nothing is hard-coded here
    traffic, issue = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        ent = tj.get({"pyramid": "k_pyramid_resize", "fast_cells": "k_fast_strips", "octree_select": "k_octree_select", "blur7": "k_blur7",
                      "orient_describe": "k_orient_describe"}[dom])
        if ent and int(ent.get("images_per_launch", -1)) == B:
            traffic = int(ent["dram_bytes_per_launch"])
            # what really bounds the kernel, from the same capture: issue slots and the busiest pipes
            issue = {k: ent[k] for k in ("issue_slots_busy_pct", "sm_throughput_pct", "alu_pipe_pct", "l1tex_throughput_pct",
                                         "lsu_shared_wavefronts_pct", "dram_throughput_pct", "warp_instructions", "ipc_active",
                                         "registers_per_thread", "achieved_occupancy_pct") if k in ent}
            if "ipc_active" in issue:
                issue["frac"] = issue["ipc_active"] / 4.0  # achieved warp-instructions per clock per SM over the 4 an SM can issue
            issue["source"] = "profiles/" + str(ent.get("report", "r2_traffic.json")) + " via profiles/r2_traffic.json (ncu --set full, one launch)"
    except Exception:
        pass
    kernel_names = {"pyramid": "k_pyramid_resize", "fast_cells": "k_fast_strips", "octree_select": "k_octree_select",
                    "blur7": "k_blur7", "orient_describe": "k_orient_describe"}
    roofline = {
        "kernel": kernel_names.get(dom, "k_" + dom), "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
        "frac": achieved / hbm_peak, "traffic": traffic, "peak_source": peak_src,
        "issue": issue,
        "bound_note": "reported against HBM as the contract asks; the ncu capture (profiles/r2_extract_kernels_full_details.txt, figures in "
                      "`issue`) shows the kernel bound by the shared-memory pipe / issue slots, DRAM at a few percent",
        "launches_per_step": dom_launches, "ms_per_step": stage_ms[dom], "algorithmic_bytes_per_step": dom_bytes,
        "stage_ms_per_step": stage_ms,
        "path": {"algorithmic_bytes_per_image": abytes_img, "achieved": value / world * abytes_img / 1e9,
                 "frac": value / world * abytes_img / 1e9 / hbm_peak, "unit": "GB/s"},
    }
    if dom_bytes_note:
        roofline["note"] = dom_bytes_note

    # ---- e2e: host buffers through the C-ABI, copies inside the timed region
    kps_h = torch.empty((B, cap, 7), dtype=torch.float32).pin_memory()
    desc_h = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    n_h = torch.zeros((B,), dtype=torch.int32).pin_memory()
    uR_h = torch.empty((P, cap), dtype=torch.float32).pin_memory()
    dep_h = torch.empty((P, cap), dtype=torch.float32).pin_memory()
    nv_h = torch.zeros((P,), dtype=torch.int32).pin_memory()
    import ctypes as C
    ptrs = (C.c_void_p * B)(*[host_imgs.data_ptr() + b * w * h for b in range(B)])
    L = orbb200.lib()

    def step_host():
        # the stereo Frame constructor's feature work as one call: extraction of L and R + ComputeStereoMatches
        orbb200.check(L.orbb200_extract_stereo_batch(ex.h, ptrs, P, w, h, w, kps_h.data_ptr(), desc_h.data_ptr(), cap, n_h.data_ptr(),
                                                     K["bf"], bconst, uR_h.data_ptr(), dep_h.data_ptr(), nv_h.data_ptr()))

    # second set of pinned result buffers: the begin / wait form keeps two steps in flight (a sequence streamed through one handle)
    kps_h2, desc_h2, n_h2 = torch.empty_like(kps_h).pin_memory(), torch.empty_like(desc_h).pin_memory(), torch.zeros_like(n_h).pin_memory()
    uR_h2, dep_h2, nv_h2 = torch.empty_like(uR_h).pin_memory(), torch.empty_like(dep_h).pin_memory(), torch.zeros_like(nv_h).pin_memory()
    outs = [(kps_h, desc_h, n_h, uR_h, dep_h, nv_h), (kps_h2, desc_h2, n_h2, uR_h2, dep_h2, nv_h2)]

    def begin_host(k):
        o = outs[k & 1]
        t = C.c_int(-1)
        orbb200.check(L.orbb200_extract_stereo_batch_begin(ex.h, ptrs, P, w, h, w, o[0].data_ptr(), o[1].data_ptr(), cap, o[2].data_ptr(),
                                                           K["bf"], bconst, o[3].data_ptr(), o[4].data_ptr(), o[5].data_ptr(), C.byref(t)))
        return t.value

    for _ in range(2):
        step_host()
    e2e_steps = max(1, min(args.steps, 10))
    # (a) one synchronous call per step
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_host()
    ex.sync()
    sync_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
    assert np.array_equal(n_h.numpy(), n_host), "host-API and device-API keypoint counts differ"
    # (b) the same steps streamed: begin(k), then wait(k - 1) -- every step's input copy and result download inside the timed region,
    #     the copies of step k under the tail of step k - 1
    for k in range(2):
        orbb200.check(L.orbb200_extract_batch_wait(ex.h, begin_host(k)))
    barrier()
    e0.record(stream)
    t0 = time.perf_counter()
    prev = -1
    for k in range(e2e_steps):
        cur = begin_host(k)
        orbb200.check(L.orbb200_extract_batch_wait(ex.h, prev))
        prev = cur
    orbb200.check(L.orbb200_extract_batch_wait(ex.h, prev))
    e1.record(stream)
    ex.sync()
    wall = time.perf_counter() - t0
    barrier()
    e2e_ms = max_over_ranks(max(e0.elapsed_time(e1), wall * 1e3))
    e2e_value = world * B * e2e_steps / (e2e_ms * 1e-3)
    e2e_sync_value = world * B * e2e_steps / (sync_ms * 1e-3)
    for o in outs:
        assert np.array_equal(o[2].numpy(), n_host) and np.array_equal(o[5].numpy(), nv_h.numpy()), "streamed steps differ from the synchronous call"
    h2d = B * w * h
    d2h = B * cap * (28 + 32) + B * 4 + 2 * P * cap * 4 + P * 4
    # the copy-bound ceiling of this step on this host: bare pinned H2D of the step's input with the D2H of its outputs running
    # beside it on a second stream, all ranks at once -- e2e can be read as a fraction of it
    d2h_d = torch.empty(d2h, dtype=torch.uint8, device=dev)
    d2h_h = torch.empty(d2h, dtype=torch.uint8).pin_memory()
    s2 = torch.cuda.Stream(device=dev)
    for _ in range(2):
        dev_imgs.copy_(host_imgs, non_blocking=True)
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    for _ in range(5):
        dev_imgs.copy_(host_imgs, non_blocking=True)
        with torch.cuda.stream(s2):
            d2h_h.copy_(d2h_d, non_blocking=True)
    torch.cuda.synchronize()
    copy_s = max_over_ranks(time.perf_counter() - t0) / 5
    barrier()
    del d2h_d, d2h_h

    # ---- config 5: brute-force top-2, database sharded by rows over the ranks
    match = bench_match(args, rank, world, dev, barrier, max_over_ranks)

    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
        "data": "synthetic",
        "config": make_config(P),
        "workload_stats": {"mean_keypoints_per_image": mean_kp, "mean_stereo_matches_per_pair": float(nv_host.mean())},
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "steps": e2e_steps,
                "api": "orbb200_extract_stereo_batch_begin / orbb200_extract_batch_wait (pinned host buffers; extraction L+R + ComputeStereoMatches in one "
                       "pipelined call per step, two steps in flight on the handle; every step's H2D and D2H inside the timed region)",
                "synchronous_call_value": e2e_sync_value,
                "synchronous_call_note": "orbb200_extract_stereo_batch, one blocking call per step: its first copy and last chunk's kernels are not overlapped",
                "h2d_gbs_per_rank": h2d / copy_s / 1e9, "copy_bound_ceiling_frames_per_s": world * B / copy_s,
                "frac_of_copy_bound": e2e_value / (world * B / copy_s), "host_bound_cpus": bound_cpus,
                "copy_ceiling_note": "bare pinned H2D of the step's input with the D2H of its outputs on a second stream, all ranks at once"},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": roofline,
        "clocks": sampler.summary(),
        "match": match,
    }
    if rank == 0 and world == 1 and not args.no_extras:
        # BASELINE configs 3 and 4 (device-resident batches) and config 1 as written (ONE frame through the C++ facade)
        del dev_imgs, kps_d, desc_d, uR_d, dep_d
        line["extra"] = {"config3_euroc_752x480_1200feat_batch256": bench_config(torch, dev, 752, 480, 1200, 8, 256, steps=5),
                         "config4_4k_3840x2160_8000feat_12lvl_batch16": bench_config(torch, dev, 3840, 2160, 8000, 12, 16, steps=5)}
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import latency_probe
            line["latency"] = latency_probe.run(iters=200, cpu_iters=5 if not args.no_cpu else 0)
        except Exception as ex_:  # the facade driver is a prebuilt binary: report, do not fail the bench line
            line["latency"] = {"error": str(ex_)[:200]}
    if rank == 0 and world == 1 and not args.no_cpu:
        import orbslam_ref
        cores = os.cpu_count() or 1
        have_ref = orbslam_ref.available()
        nthreads = max(1, cores // 2) if have_ref else cores
        sample_pairs = min(P, max(nthreads, 4) * 16)  # ~15-20 s of CPU work spread over all cores (about 1.2 s wall on 16 cores)
        fn = cpu_reference_pairs if have_ref else cpu_extract_pairs
        dt, nimg = fn(host_imgs.numpy()[:2 * sample_pairs], nthreads)
        line["cpu_baseline"] = {"value": nimg / dt, "unit": "frames/s", "cores": cores, "kind": "reference" if have_ref else "port",
                                "sample": "%d of the step's stereo pairs; %d sessions x 2 extraction threads; the reference's own "
                                          "ORBextractor.cc + stereo Frame constructor (oracle/_ref, OpenCV calls through the plain-C stand-in), "
                                          "%.1f s wall" % (sample_pairs, nthreads, dt)}
        # the variants BASELINE.md 2 lists, beside the headline baseline
        dtp, nimgp = cpu_extract_pairs(host_imgs.numpy()[:2 * sample_pairs], cores)
        line["cpu_baseline"]["variants"] = {
            "plain_c_port_all_cores": {"value": nimgp / dtp, "unit": "frames/s", "cores": cores,
                                       "note": "oracle/orb_ref.c (-O3 -march=native), one extractor pair per thread"},
            "opencv_library_view_one_thread": cpu_library_view(np.ascontiguousarray(host_imgs.numpy()[0])),
        }
        line["match"]["cpu_baseline"] = cpu_match_baseline(cores)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def bench_config(torch, dev, w, h, nfeat, nlevels, batch, steps=5):
    """device-resident extraction of `batch` same-shape synthetic frames (no stereo stage): frames/s, per-stage CUDA-event times and
    the path's HBM-roofline fraction for another BASELINE config"""
    import orbb200
    import synth
    ex = orbb200.ORBextractor(nfeat, 1.2, nlevels, 20, 7, device=dev.index)
    nbase = min(batch, 4)
    base = [synth.textured_frame(w, h, 3000 + k) for k in range(nbase)]
    host = np.empty((batch, h, w), np.uint8)
    for i in range(batch):
        host[i] = np.roll(base[i % nbase], (i // nbase) * 13, axis=0)
    imgs = torch.from_numpy(host).to(dev)
    ex._fit(w, h)
    cap = ex.cap
    kps = torch.empty((batch, cap, 7), dtype=torch.float32, device=dev)
    desc = torch.empty((batch, cap, 32), dtype=torch.uint8, device=dev)
    n = torch.zeros((batch,), dtype=torch.int32, device=dev)
    stream = torch.cuda.ExternalStream(ex.stream(), device=dev)

    def step():
        ex.extract_batch_device(imgs.data_ptr(), batch, w, h, w, w * h, kps.data_ptr(), desc.data_ptr(), cap, n.data_ptr())

    for _ in range(3):
        step()
    ex.sync()
    ex.set_profiling(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        step()
    e1.record(stream)
    ex.sync()
    ms = e0.elapsed_time(e1) / steps
    ssum, nc = ex.stage_ms_sum()
    ex.set_profiling(False)
    mean_kp = float(n.cpu().numpy().mean())
    _, abytes_img = algorithmic_bytes(w, h, 1.2, nlevels, mean_kp)
    fps = batch / (ms * 1e-3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    out = {"value": fps, "unit": "frames/s", "ms_per_step": ms, "batch": batch, "steps": steps, "mean_keypoints_per_image": mean_kp,
           "stage_ms_per_step": {k: v / max(nc, 1) for k, v in ssum.items()},
           "path_roofline": {"algorithmic_bytes_per_image": abytes_img, "achieved_gbs": fps * abytes_img / 1e9, "frac": fps * abytes_img / 1e9 / hbm}}
    ex.close()
    del imgs, kps, desc
    torch.cuda.empty_cache()
    return out


def bench_match(args, rank, world, dev, barrier, max_over_ranks):
    import torch
    import torch.distributed as dist
    import orbb200
    import synth
    L = orbb200.lib()
    q, db, _ = synth.descriptor_db(NQ, NDB, seed=7)
    lo, hi = orbb200.shard_range(NDB, world, rank)
    per = hi - lo
    q_d = torch.from_numpy(q).to(dev)
    db_d = torch.from_numpy(np.ascontiguousarray(db[lo:lo + per])).to(dev)
    part_d = torch.empty((NQ, 4), dtype=torch.int32, device=dev)
    out_d = torch.empty((NQ, 4), dtype=torch.int32, device=dev)
    ts = torch.cuda.Stream(device=dev)  # a real (non-default) stream: kernels, NCCL and events all order on it
    torch.cuda.set_stream(ts)
    st = ts.cuda_stream

    # N > 1: the exchange is fused into the search kernel (peer stores over NVLink + flags, one launch per step);
    # the NCCL all-gather + merge formulation is timed beside it as the baseline it replaces
    sm = orbb200.ShardedMatcher(rank, world, NQ, dev.index) if world > 1 else None

    def step_nccl():
        orbb200.check(L.orbb200_match_bruteforce_top2_device(q_d.data_ptr(), NQ, db_d.data_ptr(), per, part_d.data_ptr(), lo, st))
        if world > 1:
            gath_d = orbb200.gather_top2_parts(part_d)
            orbb200.check(L.orbb200_merge_top2_device(gath_d.data_ptr(), world, NQ, out_d.data_ptr(), st))

    def step_fused():
        sm.match_device(q_d.data_ptr(), NQ, db_d.data_ptr(), per, lo, out_d.data_ptr(), st)

    def timed(step):
        for _ in range(3):
            step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1)) / steps

    steps = max(args.steps, 5)
    ms_nccl = timed(step_nccl)
    nccl_ref = out_d.clone() if world > 1 else None
    if world > 1:
        out_d.zero_()
        ms = timed(step_fused)
        sm.status(st)
        assert torch.equal(out_d, nccl_ref), "fused exchange and NCCL all-gather + merge disagree"
    else:
        ms = ms_nccl
    pairs = float(NQ) * NDB
    popc = orbb200.popc_peak(dev.index)
    val = pairs / (ms * 1e-3)
    res = out_d if world > 1 else part_d
    r = res.cpu().numpy()
    m = {
        "metric": "Hamming match pairs/s (brute-force top-2, 2048 queries x 1048576 rows)", "value": val, "unit": "pairs/s",
        "ms_per_step": ms, "scaling": "strong",
        "sharding": "database rows / rank; exchange fused into k_bf_top2<true> (peer stores over NVLink + release/acquire flags, "
                    "no NCCL on the data path)" if world > 1 else "single GPU",
        "gpu_launches_per_step": 1 if world > 1 else 2,
        # ALGORITHMIC work = 8 popc32 per pair (32-byte rows).  The kernel executes 5 POPC + 14 LOP3 per pair (carry-save adders fold
        # seven XOR words into one word of weight 1 and three of weight 2), so `frac` against the POPC pipe can exceed 1.
        "roofline": {"kernel": "k_bf_top2", "bound": "popc", "achieved": val * 8 / world / 1e12, "unit": "Tpopc32/s (algorithmic: 8 per pair)",
                     "peak": 16 * 148 * 1.965e9 / 1e12, "peak_source": "nominal POPC rate 16 / clk / SM x 148 SMs x 1.965 GHz",
                     "frac": val * 8 / world / (16 * 148 * 1.965e9),
                     "executed_popc_per_pair": 5, "executed_frac_of_nominal": val * 5 / world / (16 * 148 * 1.965e9),
                     "measured_popc_peak": popc / 1e12, "frac_of_measured_peak": val * 8 / world / popc},
        "checksum": int(r[:, 1].astype(np.int64).sum()),
    }
    if world > 1:
        m["nccl_allgather_baseline"] = {"ms_per_step": ms_nccl, "value": pairs / (ms_nccl * 1e-3), "unit": "pairs/s",
                                        "note": "k_bf_top2 + k_merge_top2 + ncclAllGather + k_merge_top2 (3 launches + 1 collective)"}
    return m


def cpu_match_baseline(cores):
    """the DescriptorDistance loop of ORBmatcher.cc:1733-1749 (bit-hack popcount, as the reference writes it) on all host cores, with
    the __builtin_popcountll variant beside it"""
    import ref_cpu
    import synth
    from concurrent.futures import ThreadPoolExecutor
    nq, ndb = 256, 1 << 16
    q, db, _ = synth.descriptor_db(nq, ndb, seed=7)
    chunks = np.array_split(np.arange(nq), cores)
    res = {}
    with ThreadPoolExecutor(cores) as ex:
        for name, pop in (("bit_hack", False), ("popcountll", True)):
            def work(ix, pop=pop):
                if len(ix):
                    ref_cpu.bf_top2(q[ix], db, use_popcnt=pop)
            list(ex.map(work, chunks))
            t0 = time.perf_counter()
            reps = 4
            for _ in range(reps):
                list(ex.map(work, chunks))
            res[name] = reps * nq * ndb / (time.perf_counter() - t0)
    return {"value": res["bit_hack"], "unit": "pairs/s", "cores": cores, "kind": "port",
            "sample": "%d queries x %d rows x 4 reps, queries split over threads; the reference's bit-hack popcount" % (nq, ndb),
            "popcountll_variant": {"value": res["popcountll"], "unit": "pairs/s"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--pairs", type=int, default=160, help="stereo pairs per step per GPU (2x images)")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-bind", action="store_true", help="do not pin the rank to its GPU's local CPUs (A/B of the e2e feed)")
    ap.add_argument("--no-extras", action="store_true", help="skip the config 3 / config 4 / latency objects")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world == 1 and args.gpus > 1:
        # plain `python bench.py --gpus N`: re-launch under torchrun, one rank per GPU
        import subprocess
        port = 29500 + os.getpid() % 2000
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
