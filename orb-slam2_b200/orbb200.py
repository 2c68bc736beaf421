"""ctypes binding of liborbb200.so (include/orbb200.h) for the tests and bench.py.

The host-side product is C++ (include/orbb200/ORBextractor.h, ORBmatcher.h); this module only lets the
Python harness drive the same C-ABI.  Names mirror the reference classes (ORBextractor, ORBmatcher).
There is NO fallback: if the shared library is missing or CUDA is unavailable every call raises.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ORBB200_LIB_PATH", os.path.join(HERE, "liborbb200.so"))  # override: A/B builds of the library in tuning runs

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
TOP2_DTYPE = np.dtype([("best_dist", "<i4"), ("best_idx", "<i4"), ("second_dist", "<i4"), ("second_idx", "<i4")])

EXPORTS = [
    "orbb200_extractor_create", "orbb200_extractor_destroy", "orbb200_level_tables", "orbb200_compute_level_tables", "orbb200_max_keypoints", "orbb200_max_keypoints_for_shape", "orbb200_last_capacity", "orbb200_pyramid_layout", "orbb200_pyramid_download", "orbb200_pyramid_map", "orbb200_host_alloc", "orbb200_host_free", "orbb200_window_search_resident", "orbb200_extractor_outputs_device", "orbb200_device_alloc", "orbb200_device_free", "orbb200_copy_to_device", "orbb200_copy_to_host", "orbb200_copy_on_device", "orbb200_h2d_bytes_total", "orbb200_device_sync", "orbb200_bind_host_to_device", "orbb200_extractor_set_graph", "orbb200_extractor_set_octree_depth", "orbb200_remap_create", "orbb200_remap_destroy", "orbb200_remap_batch", "orbb200_remap_batch_device", "orbb200_init_undistort_rectify_map", "orbb200_extract_batch_begin", "orbb200_extract_stereo_batch_begin", "orbb200_extract_batch_wait",
    "orbb200_extract", "orbb200_extract_batch", "orbb200_extract_batch_device", "orbb200_pyramid_level",
    "orbb200_extractor_stream", "orbb200_extractor_sync", "orbb200_extractor_set_profiling", "orbb200_extractor_set_pipeline",
    "orbb200_extractor_stage_ms", "orbb200_extractor_stage_ms_sum", "orbb200_extractor_last_launches", "orbb200_debug_candidates",
    "orbb200_debug_blurred_level", "orbb200_descriptor_distance", "orbb200_match_bruteforce_top2",
    "orbb200_match_bruteforce_top2_device", "orbb200_match_bruteforce_top2_device_noalloc", "orbb200_match_candidates", "orbb200_candidate_distances", "orbb200_search_triangulation", "orbb200_window_search", "orbb200_merge_top2",
    "orbb200_merge_top2_device", "orbb200_stereo_match", "orbb200_stereo_match_batch",
    "orbb200_stereo_match_batch_device", "orbb200_extract_stereo_batch", "orbb200_popc_peak", "orbb200_last_error", "orbb200_version",
    "orbb200_vocabulary_load_text", "orbb200_vocabulary_create", "orbb200_vocabulary_destroy", "orbb200_vocabulary_info",
    "orbb200_vocabulary_transform", "orbb200_vocabulary_transform_device", "orbb200_compute_bow_batch",
    "orbb200_compute_bow_batch_device", "orbb200_distinctive_descriptors", "orbb200_distinctive_descriptors_device",
    "orbb200_exchange_create", "orbb200_exchange_connect", "orbb200_exchange_destroy", "orbb200_match_sharded_top2_device",
    "orbb200_exchange_status", "orbb200_extract_color", "orbb200_extract_color_batch", "orbb200_cvt_gray",
    "orbb200_undistort_keypoints", "orbb200_undistort_keypoints_device", "orbb200_window_search_device",
]

_lib = None


class OrbB200Error(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OrbB200Error("liborbb200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                           "there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, ip, fp, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    L.orbb200_last_error.restype = C.c_char_p
    L.orbb200_version.restype = C.c_char_p
    L.orbb200_extractor_create.argtypes = [C.POINTER(vp), ip, fp, ip, ip, ip, ip]
    L.orbb200_extractor_destroy.argtypes = [vp]
    L.orbb200_extractor_destroy.restype = None
    L.orbb200_level_tables.argtypes = [vp, vp, vp, vp, vp, vp]
    L.orbb200_compute_level_tables.argtypes = [ip, fp, ip, vp, vp, vp, vp, vp, vp]
    L.orbb200_max_keypoints.argtypes = [vp]
    L.orbb200_max_keypoints_for_shape.argtypes = [vp, ip, ip]
    L.orbb200_last_capacity.argtypes = [vp]
    L.orbb200_pyramid_layout.argtypes = [vp, vp, vp, vp, vp, vp]
    L.orbb200_pyramid_download.argtypes = [vp, ip, vp, sz]
    L.orbb200_pyramid_map.argtypes = [vp, ip, vp]
    L.orbb200_extractor_set_octree_depth.argtypes = [vp, ip]
    L.orbb200_extract_batch_begin.argtypes = [vp, vp, ip, ip, ip, ip, vp, vp, ip, vp, vp]
    L.orbb200_extract_stereo_batch_begin.argtypes = [vp, vp, ip, ip, ip, ip, vp, vp, ip, vp, fp, fp, vp, vp, vp, vp]
    L.orbb200_extract_batch_wait.argtypes = [vp, ip]
    L.orbb200_remap_create.argtypes = [vp, vp, vp, ip, ip, ip]
    L.orbb200_remap_destroy.argtypes = [vp]
    L.orbb200_remap_destroy.restype = None
    L.orbb200_remap_batch.argtypes = [vp, vp, ip, ip, ip, ip, vp, ip]
    L.orbb200_remap_batch_device.argtypes = [vp, vp, ip, ip, ip, ip, C.c_size_t, vp, ip, C.c_size_t, vp]
    L.orbb200_init_undistort_rectify_map.argtypes = [vp, vp, ip, vp, vp, ip, ip, vp, vp]
    L.orbb200_extractor_set_graph.argtypes = [vp, ip]
    L.orbb200_extract.argtypes = [vp, vp, ip, ip, ip, vp, vp, ip, vp]
    L.orbb200_extract_batch.argtypes = [vp, vp, ip, ip, ip, ip, vp, vp, ip, vp]
    L.orbb200_extract_batch_device.argtypes = [vp, vp, ip, ip, ip, ip, sz, vp, vp, ip, vp]
    L.orbb200_pyramid_level.argtypes = [vp, ip, ip, vp, ip, vp, vp]
    L.orbb200_extractor_stream.argtypes = [vp]
    L.orbb200_extractor_stream.restype = vp
    L.orbb200_extractor_sync.argtypes = [vp]
    L.orbb200_extractor_set_profiling.argtypes = [vp, ip]
    L.orbb200_extractor_set_pipeline.argtypes = [vp, ip, ip, ip]
    L.orbb200_extractor_stage_ms.argtypes = [vp, vp, ip]
    L.orbb200_extractor_stage_ms_sum.argtypes = [vp, vp, ip, vp]
    L.orbb200_extractor_last_launches.argtypes = [vp]
    L.orbb200_debug_candidates.argtypes = [vp, ip, ip, vp, ip, vp]
    L.orbb200_debug_blurred_level.argtypes = [vp, ip, ip, vp, ip]
    L.orbb200_descriptor_distance.argtypes = [vp, vp, ip, vp, ip]
    L.orbb200_match_bruteforce_top2.argtypes = [vp, ip, vp, ip, vp, ip]
    L.orbb200_match_bruteforce_top2_device.argtypes = [vp, ip, vp, ip, vp, ip, vp]
    L.orbb200_match_bruteforce_top2_device_noalloc.argtypes = [vp, ip, vp, ip, vp, ip, vp]
    L.orbb200_match_candidates.argtypes = [vp, ip, vp, ip, vp, vp, vp, ip]
    L.orbb200_candidate_distances.argtypes = [vp, ip, vp, ip, vp, vp, vp, ip]
    L.orbb200_search_triangulation.argtypes = [vp, vp, vp, ip, vp, vp, vp, ip, vp, vp, vp, fp, fp, vp, vp, ip, ip, vp, vp, ip]
    L.orbb200_window_search.argtypes = [vp, vp, ip, fp, fp, fp, fp, ip, ip, vp, vp, ip, vp, vp, vp, ip, ip]
    L.orbb200_merge_top2.argtypes = [vp, ip, ip, vp, ip]
    L.orbb200_merge_top2_device.argtypes = [vp, ip, ip, vp, vp]
    L.orbb200_distinctive_descriptors.argtypes = [vp, vp, ip, vp, vp, ip]
    L.orbb200_extract_color.argtypes = [vp, vp, ip, ip, ip, ip, ip, vp, vp, ip, vp]
    L.orbb200_extract_color_batch.argtypes = [vp, vp, ip, ip, ip, ip, ip, ip, vp, vp, ip, vp]
    L.orbb200_cvt_gray.argtypes = [vp, ip, ip, ip, ip, ip, vp, ip, ip]
    L.orbb200_undistort_keypoints.argtypes = [vp, ip, vp, vp, vp, ip]
    L.orbb200_undistort_keypoints_device.argtypes = [vp, ip, vp, vp, vp, vp]
    L.orbb200_window_search_device.argtypes = [vp, vp, ip, vp, fp, fp, fp, fp, ip, ip, vp, vp, ip, vp, vp, vp, ip, vp, vp]
    L.orbb200_exchange_create.argtypes = [vp, ip, ip, ip, ip, vp]
    L.orbb200_exchange_connect.argtypes = [vp, vp]
    L.orbb200_exchange_destroy.argtypes = [vp]
    L.orbb200_exchange_destroy.restype = None
    L.orbb200_match_sharded_top2_device.argtypes = [vp, vp, ip, vp, ip, ip, vp, vp]
    L.orbb200_exchange_status.argtypes = [vp, vp]
    L.orbb200_distinctive_descriptors_device.argtypes = [vp, vp, ip, vp, vp, vp]
    L.orbb200_stereo_match.argtypes = [vp, vp, fp, fp, vp, vp, vp]
    L.orbb200_stereo_match_batch.argtypes = [vp, ip, fp, fp, vp, vp, vp]
    L.orbb200_stereo_match_batch_device.argtypes = [vp, ip, fp, fp, vp, vp, vp]
    L.orbb200_extract_stereo_batch.argtypes = [vp, vp, ip, ip, ip, ip, vp, vp, ip, vp, fp, fp, vp, vp, vp]
    L.orbb200_popc_peak.argtypes = [ip, C.POINTER(C.c_double)]
    L.orbb200_vocabulary_load_text.argtypes = [C.POINTER(vp), C.c_char_p, ip]
    L.orbb200_vocabulary_create.argtypes = [C.POINTER(vp), ip, ip, ip, ip, ip, vp, vp, vp, vp, ip]
    L.orbb200_vocabulary_destroy.argtypes = [vp]
    L.orbb200_vocabulary_destroy.restype = None
    L.orbb200_vocabulary_info.argtypes = [vp, vp, vp, vp, vp, vp, vp]
    L.orbb200_vocabulary_transform.argtypes = [vp, vp, ip, ip, vp, vp, vp]
    L.orbb200_vocabulary_transform_device.argtypes = [vp, vp, ip, ip, vp, vp, vp, vp]
    L.orbb200_compute_bow_batch.argtypes = [vp, vp, vp, ip, ip, vp, vp, vp, vp, vp, vp, vp]
    L.orbb200_compute_bow_batch_device.argtypes = [vp, vp, vp, ip, ip, ip, vp, vp, vp, vp, vp, vp]
    _lib = L
    return L


def popc_peak(device=-1):
    v = C.c_double()
    check(lib().orbb200_popc_peak(device, C.byref(v)))
    return v.value


def check(rc):
    if rc != 0:
        raise OrbB200Error("orbb200 error %d: %s" % (rc, lib().orbb200_last_error().decode()))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class ORBextractor:
    """Same constructor and accessors as ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:66-100)."""

    STAGES = ("pyramid", "fast_cells", "octree_select", "blur7", "orient_describe")

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=-1):
        self.L = lib()
        h = C.c_void_p()
        check(self.L.orbb200_extractor_create(C.byref(h), nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device))
        self.h = h
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self.cap = self.L.orbb200_max_keypoints(self.h)
        n = nlevels
        self._t = [np.zeros(n, np.float32) for _ in range(4)] + [np.zeros(n, np.int32)]
        check(self.L.orbb200_level_tables(self.h, *[_p(a) for a in self._t]))

    def close(self):
        if getattr(self, "h", None):
            self.L.orbb200_extractor_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _fit(self, w, h):
        """grow self.cap to the bound for this image shape: max(nfeat + 3, 4 * nIni) per level (a wide image with a small quota
        returns more keypoints than its quota, src/ORBextractor.cc:624-760)"""
        self.cap = max(self.cap, self.L.orbb200_max_keypoints_for_shape(self.h, int(w), int(h)))

    # accessors (reference ORBextractor.h:78-98)
    def GetLevels(self):
        return self.nlevels

    def GetScaleFactors(self):
        return self._t[0].copy()

    def GetInverseScaleFactors(self):
        return self._t[1].copy()

    def GetScaleSigmaSquares(self):
        return self._t[2].copy()

    def GetInverseScaleSigmaSquares(self):
        return self._t[3].copy()

    def GetFeaturesPerLevel(self):
        return self._t[4].copy()

    def __call__(self, image):
        """operator()(image, mask, keypoints, descriptors): returns (keypoints[KP_DTYPE], descriptors[N,32])."""
        if image is None or image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "CV_8UC1 required (reference :1269)"
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        self._fit(w, h)
        kps = np.zeros(self.cap, KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = C.c_int(0)
        check(self.L.orbb200_extract(self.h, _p(image), w, h, image.strides[0], _p(kps), _p(desc), self.cap, C.byref(n)))
        return kps[:n.value].copy(), desc[:n.value].copy()

    def extract_color(self, image, r_first):
        """Tracking::GrabImage* prologue + operator(): image [H, W, 3|4] uint8, r_first = mbRGB.  The gray conversion runs
        on the device inside the level-0 staging; mImGray is pyramid_level(0)."""
        assert image.dtype == np.uint8 and image.ndim == 3 and image.shape[2] in (3, 4)
        if image.strides[2] != 1 or image.strides[1] != image.shape[2]:
            image = np.ascontiguousarray(image)
        h, w, ch = image.shape
        self._fit(w, h)
        kps = np.zeros(self.cap, KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = C.c_int(0)
        check(self.L.orbb200_extract_color(self.h, _p(image), w, h, image.strides[0], ch, int(bool(r_first)), _p(kps), _p(desc),
                                           self.cap, C.byref(n)))
        return kps[:n.value].copy(), desc[:n.value].copy()

    def extract_color_batch(self, images, r_first):
        """images: [B, H, W, 3|4] uint8 (host).  Returns (kps[B,cap], desc[B,cap,32], n[B])."""
        B, h, w, ch = images.shape
        self._fit(w, h)
        assert images.dtype == np.uint8 and images.strides[3] == 1 and images.strides[2] == ch
        ptrs = (C.c_void_p * B)(*[images.ctypes.data + b * images.strides[0] for b in range(B)])
        kps = np.zeros((B, self.cap), KP_DTYPE)
        desc = np.zeros((B, self.cap, 32), np.uint8)
        n_out = np.zeros(B, np.int32)
        check(self.L.orbb200_extract_color_batch(self.h, ptrs, B, w, h, images.strides[1], ch, int(bool(r_first)), _p(kps), _p(desc),
                                                 self.cap, _p(n_out)))
        return kps, desc, n_out

    def extract_batch(self, images, kps=None, desc=None, n_out=None):
        """images: [B,H,W] uint8 (host).  Returns (kps[B,cap], desc[B,cap,32], n[B])."""
        images = np.ascontiguousarray(images, np.uint8) if not isinstance(images, np.ndarray) else images
        B, h, w = images.shape
        self._fit(w, h)
        stride = images.strides[1]
        ptrs = (C.c_void_p * B)(*[images.ctypes.data + b * images.strides[0] for b in range(B)])
        if kps is None:
            kps = np.zeros((B, self.cap), KP_DTYPE)
            desc = np.zeros((B, self.cap, 32), np.uint8)
            n_out = np.zeros(B, np.int32)
        check(self.L.orbb200_extract_batch(self.h, ptrs, B, w, h, stride, _p(kps), _p(desc), kps.shape[1], _p(n_out)))
        return kps, desc, n_out

    def extract_stereo_batch(self, images, bf, b):
        """images: [2P,H,W] uint8 (host), pair p = (2p, 2p+1).  Extraction + ComputeStereoMatches in one pipelined call.
        Returns (kps, desc, n, uRight[P,cap], depth[P,cap], n_valid[P])."""
        B, h, w = images.shape
        self._fit(w, h)
        ptrs = (C.c_void_p * B)(*[images.ctypes.data + i * images.strides[0] for i in range(B)])
        kps = np.zeros((B, self.cap), KP_DTYPE)
        desc = np.zeros((B, self.cap, 32), np.uint8)
        n_out = np.zeros(B, np.int32)
        uR = np.zeros((B // 2, self.cap), np.float32)
        depth = np.zeros((B // 2, self.cap), np.float32)
        nv = np.zeros(B // 2, np.int32)
        check(self.L.orbb200_extract_stereo_batch(self.h, ptrs, B // 2, w, h, images.strides[1], _p(kps), _p(desc), self.cap, _p(n_out),
                                                  bf, b, _p(uR), _p(depth), _p(nv)))
        return kps, desc, n_out, uR, depth, nv

    def extract_batch_device(self, imgs_ptr, batch, width, height, stride, image_stride, kps_ptr, desc_ptr, cap, n_ptr):
        check(self.L.orbb200_extract_batch_device(self.h, imgs_ptr, batch, width, height, stride, image_stride,
                                                  kps_ptr, desc_ptr, cap, n_ptr))

    def stereo_match_batch(self, npairs, bf, b):
        """Frame::ComputeStereoMatches for the pairs (image 2p, 2p+1) of the last extract_batch call (host outputs)."""
        uR = np.zeros((npairs, self.cap), np.float32)
        depth = np.zeros((npairs, self.cap), np.float32)
        nv = np.zeros(npairs, np.int32)
        check(self.L.orbb200_stereo_match_batch(self.h, npairs, bf, b, _p(uR), _p(depth), _p(nv)))
        return uR, depth, nv

    def stereo_match_batch_device(self, npairs, bf, b, uR_ptr, depth_ptr, nv_ptr):
        check(self.L.orbb200_stereo_match_batch_device(self.h, npairs, bf, b, uR_ptr, depth_ptr, nv_ptr))

    def stream(self):
        return self.L.orbb200_extractor_stream(self.h)

    def sync(self):
        check(self.L.orbb200_extractor_sync(self.h))

    def set_pipeline(self, chunk_images, compute_streams, device_chunk_images=0):
        check(self.L.orbb200_extractor_set_pipeline(self.h, int(chunk_images), int(compute_streams), int(device_chunk_images)))

    def set_octree_depth(self, max_depth):
        check(self.L.orbb200_extractor_set_octree_depth(self.h, int(max_depth)))

    def set_graph(self, on):
        check(self.L.orbb200_extractor_set_graph(self.h, int(on)))

    def pyramid(self, image_index=0):
        """all un-blurred levels of one image of the last call from ONE device-to-host copy of its slab"""
        n = self.nlevels
        offs, pit, ws, hs = (np.zeros(n, np.int32) for _ in range(4))
        nb = C.c_size_t(0)
        check(self.L.orbb200_pyramid_layout(self.h, _p(offs), _p(pit), _p(ws), _p(hs), C.byref(nb)))
        slab = np.zeros(nb.value, np.uint8)
        check(self.L.orbb200_pyramid_download(self.h, image_index, _p(slab), nb.value))
        return [slab[offs[l]:offs[l] + pit[l] * hs[l]].reshape(hs[l], pit[l])[:, :ws[l]] for l in range(n)]

    def set_profiling(self, on):
        check(self.L.orbb200_extractor_set_profiling(self.h, int(on)))

    def stage_ms(self):
        ms = np.zeros(5, np.float32)
        check(self.L.orbb200_extractor_stage_ms(self.h, _p(ms), 5))
        return dict(zip(self.STAGES, ms.tolist()))

    def stage_ms_sum(self):
        """(dict of summed stage ms, number of calls covered) since set_profiling(True)."""
        ms = np.zeros(6, np.float32)
        n = C.c_int()
        check(self.L.orbb200_extractor_stage_ms_sum(self.h, _p(ms), 6, C.byref(n)))
        self.extract_total_ms_sum = float(ms[5])  # first launch to last (the stages run back to back on one stream)
        return dict(zip(self.STAGES, ms[:5].tolist())), n.value

    def last_launches(self):
        return self.L.orbb200_extractor_last_launches(self.h)

    def pyramid_level(self, level, image_index=0):
        """mvImagePyramid[level] (un-blurred) of the last call."""
        w, h = C.c_int(), C.c_int()
        check(self.L.orbb200_pyramid_level(self.h, image_index, level, None, 0, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), np.uint8)
        check(self.L.orbb200_pyramid_level(self.h, image_index, level, _p(out), w.value, C.byref(w), C.byref(h)))
        return out

    def debug_blurred_level(self, level, image_index=0):
        w, h = C.c_int(), C.c_int()
        check(self.L.orbb200_pyramid_level(self.h, image_index, level, None, 0, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), np.uint8)
        check(self.L.orbb200_debug_blurred_level(self.h, image_index, level, _p(out), w.value))
        return out

    def debug_candidates(self, level, image_index=0):
        n = C.c_int()
        check(self.L.orbb200_debug_candidates(self.h, image_index, level, None, 0, C.byref(n)))
        out = np.zeros((max(n.value, 1), 3), np.int32)
        check(self.L.orbb200_debug_candidates(self.h, image_index, level, _p(out), n.value, C.byref(n)))
        return out[:n.value]


class ORBmatcher:
    """Descriptor-distance searches of ORB_SLAM2::ORBmatcher (reference include/ORBmatcher.h:38-100) on POD arrays."""
    TH_HIGH = 100
    TH_LOW = 50
    HISTO_LENGTH = 30

    def __init__(self, nnratio=0.6, checkOri=True, device=-1):
        self.L = lib()
        self.mfNNratio = np.float32(nnratio)
        self.mbCheckOrientation = checkOri
        self.device = device

    def DescriptorDistance(self, a, b):
        a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
        b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
        out = np.zeros(len(a), np.int32)
        check(self.L.orbb200_descriptor_distance(_p(a), _p(b), len(a), _p(out), self.device))
        return out

    def bruteforce_top2(self, q, db):
        q = np.ascontiguousarray(q, np.uint8)
        db = np.ascontiguousarray(db, np.uint8)
        out = np.zeros(len(q), TOP2_DTYPE)
        check(self.L.orbb200_match_bruteforce_top2(_p(q), len(q), _p(db), len(db), _p(out), self.device))
        return out

    def match_candidates(self, q, t, offsets, idx):
        q = np.ascontiguousarray(q, np.uint8)
        t = np.ascontiguousarray(t, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.int32)
        idx = np.ascontiguousarray(idx, np.int32)
        out = np.zeros(len(q), TOP2_DTYPE)
        check(self.L.orbb200_match_candidates(_p(q), len(q), _p(t), len(t), _p(offsets), _p(idx), _p(out), self.device))
        return out

    def candidate_distances(self, q, t, offsets, idx):
        q = np.ascontiguousarray(q, np.uint8)
        t = np.ascontiguousarray(t, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.int32)
        idx = np.ascontiguousarray(idx, np.int32)
        out = np.zeros(max(int(offsets[-1]), 1), np.uint16)
        check(self.L.orbb200_candidate_distances(_p(q), len(q), _p(t), len(t), _p(offsets), _p(idx), _p(out), self.device))
        return out[:int(offsets[-1])]

    def search_triangulation(self, q, qxy, qstereo, t, tkp, tstereo, offs, idx, F12, ex, ey, sf, sig2, th_low=50):
        """SearchForTriangulation resolved on the device (orbb200_search_triangulation): (best_idx2, best_dist) per query"""
        q, t = np.ascontiguousarray(q, np.uint8), np.ascontiguousarray(t, np.uint8)
        qxy = np.ascontiguousarray(qxy, np.float32)
        qstereo, tstereo = np.ascontiguousarray(qstereo, np.uint8), np.ascontiguousarray(tstereo, np.uint8)
        tkp = np.ascontiguousarray(tkp, KP_DTYPE)
        offs, idx = np.ascontiguousarray(offs, np.int32), np.ascontiguousarray(idx, np.int32)
        F12, sf, sig2 = np.ascontiguousarray(F12, np.float32), np.ascontiguousarray(sf, np.float32), np.ascontiguousarray(sig2, np.float32)
        bi, bd = np.zeros(len(q), np.int32), np.zeros(len(q), np.int32)
        check(self.L.orbb200_search_triangulation(_p(q), _p(qxy), _p(qstereo), len(q), _p(t), _p(tkp), _p(tstereo), len(t), _p(offs), _p(idx), _p(F12),
                                                  ex, ey, _p(sf), _p(sig2), len(sf), th_low, _p(bi), _p(bd), self.device))
        return bi, bd

    def window_search(self, tkps, tdesc, grid, windows, qdesc, cap=None):
        """F3 + F2 + distances (orbb200_window_search).  grid = (minX, minY, invW, invH, cols, rows);
        windows [nq,5] float32 = x, y, r, minLevel, maxLevel.  Returns (offsets[nq+1], cand_idx, dist)."""
        tkps = np.ascontiguousarray(tkps, KP_DTYPE)
        tdesc = np.ascontiguousarray(tdesc, np.uint8)
        windows = np.ascontiguousarray(windows, np.float32)
        qdesc = np.ascontiguousarray(qdesc, np.uint8)
        nq = len(windows)
        offs = np.zeros(nq + 1, np.int32)
        cap = cap if cap is not None else max(64 * nq, 1)
        minX, minY, invW, invH, cols, rows = grid
        while True:
            idx = np.zeros(max(cap, 1), np.int32)
            dist = np.zeros(max(cap, 1), np.uint16)
            rc = self.L.orbb200_window_search(_p(tkps), _p(tdesc), len(tkps), minX, minY, invW, invH, cols, rows, _p(windows),
                                              _p(qdesc), nq, _p(offs), _p(idx), _p(dist), cap, self.device)
            if rc == -3 and offs[-1] > cap:  # ORBB200_ERR_CAPACITY: offsets are valid, retry with the exact size
                cap = int(offs[-1])
                continue
            check(rc)
            return offs, idx[:offs[-1]], dist[:offs[-1]]

    def merge_top2(self, parts):
        parts = np.ascontiguousarray(parts, TOP2_DTYPE)
        ns, nq = parts.shape
        out = np.zeros(nq, TOP2_DTYPE)
        check(self.L.orbb200_merge_top2(_p(parts), ns, nq, _p(out), self.device))
        return out

    def distinctive_descriptors(self, desc, offsets):
        """MapPoint::ComputeDistinctiveDescriptors (reference MapPoint.cc:266-340) for a batch of map points:
        desc [total, 32] u8, offsets [G+1] -> (best row per group relative to the group, its median distance)."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        offsets = np.ascontiguousarray(offsets, np.int32)
        g = len(offsets) - 1
        idx = np.zeros(max(g, 0), np.int32)
        med = np.zeros(max(g, 0), np.int32)
        check(self.L.orbb200_distinctive_descriptors(_p(desc), _p(offsets), g, _p(idx), _p(med), self.device))
        return idx, med

    def accept_ratio(self, top2, th=None):
        """best <= th and best < ratio*second (reference ORBmatcher.cc:242-244), f32 compare."""
        th = self.TH_LOW if th is None else th
        b = top2["best_dist"].astype(np.float32)
        s = top2["second_dist"].astype(np.float32)
        return (top2["best_dist"] <= th) & (b < self.mfNNratio * s)


class ORBVocabulary:
    """ORB_SLAM2::ORBVocabulary (include/ORBVocabulary.h:31-32 = DBoW2::TemplatedVocabulary<cv::Mat, FORB>): the tree
    lives in HBM, transform() walks a batch of descriptors on the GPU."""

    def __init__(self, device=-1):
        self.h = C.c_void_p()
        self.device = device

    def _drop(self):
        if getattr(self, "h", None) is not None and self.h.value:
            lib().orbb200_vocabulary_destroy(self.h)
            self.h = C.c_void_p()

    close = _drop

    def __del__(self):
        try:
            self._drop()
        except Exception:
            pass

    def loadFromTextFile(self, path):
        """TemplatedVocabulary::loadFromTextFile (TemplatedVocabulary.h:1358-1443); False on a bad file like the reference."""
        self._drop()
        rc = lib().orbb200_vocabulary_load_text(C.byref(self.h), path.encode(), self.device)
        if rc == -2:
            check(rc)
        return rc == 0

    @classmethod
    def from_arrays(cls, k, L, scoring, weighting, parent, is_leaf, desc, weight, device=-1):
        v = cls(device)
        parent = np.ascontiguousarray(parent, dtype=np.int32)
        is_leaf = np.ascontiguousarray(is_leaf, dtype=np.uint8)
        desc = np.ascontiguousarray(desc, dtype=np.uint8)
        weight = np.ascontiguousarray(weight, dtype=np.float64)
        check(lib().orbb200_vocabulary_create(C.byref(v.h), k, L, scoring, weighting, len(parent), _p(parent), _p(is_leaf), _p(desc),
                                              _p(weight), device))
        return v

    def info(self):
        a = np.zeros(6, np.int32)
        check(lib().orbb200_vocabulary_info(self.h, *[C.c_void_p(a.ctypes.data + 4 * i) for i in range(6)]))
        return dict(zip(("k", "L", "scoring", "weighting", "nodes", "words"), [int(x) for x in a]))

    def size(self):
        return self.info()["words"]

    def empty(self):
        return self.size() == 0

    def transform_features(self, desc, levelsup=0):
        """per-feature transform(feature, word_id, weight, nid, levelsup) (:1218-1259) for every row of desc."""
        desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        n = len(desc)
        word, weight, nid = np.zeros(n, np.int32), np.zeros(n, np.float64), np.zeros(n, np.int32)
        check(lib().orbb200_vocabulary_transform(self.h, _p(desc), n, levelsup, _p(word), _p(weight), _p(nid)))
        return word, weight, nid

    def transform_features_device(self, desc_ptr, n, levelsup, word_ptr, weight_ptr, nid_ptr, stream=None):
        check(lib().orbb200_vocabulary_transform_device(self.h, C.c_void_p(desc_ptr), n, levelsup, C.c_void_p(word_ptr),
                                                        C.c_void_p(weight_ptr), C.c_void_p(nid_ptr), C.c_void_p(stream or 0)))

    def compute_bow_batch_device(self, desc_ptr, n_ptr, nimg, cap, levelsup, bow_id_ptr, bow_val_ptr, n_bow_ptr, fv_key_ptr, n_kept_ptr,
                                 stream=None):
        check(lib().orbb200_compute_bow_batch_device(self.h, C.c_void_p(desc_ptr), C.c_void_p(n_ptr), nimg, cap, levelsup,
                                                     C.c_void_p(bow_id_ptr), C.c_void_p(bow_val_ptr), C.c_void_p(n_bow_ptr),
                                                     C.c_void_p(fv_key_ptr), C.c_void_p(n_kept_ptr), C.c_void_p(stream or 0)))

    def compute_bow_batch_raw(self, desc, counts, levelsup=4):
        desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        counts = np.ascontiguousarray(counts, dtype=np.int32)
        assert int(counts.sum()) == len(desc)
        n, m = len(desc), len(counts)
        out = dict(bow_off=np.zeros(m + 1, np.int32), bow_id=np.zeros(n + 1, np.int32), bow_val=np.zeros(n + 1, np.float64),
                   fv_off=np.zeros(m + 1, np.int32), fv_node=np.zeros(n + 1, np.int32), fv_feat_off=np.zeros(n + 2, np.int32),
                   fv_feat=np.zeros(n + 1, np.int32))
        check(lib().orbb200_compute_bow_batch(self.h, _p(desc), _p(counts), m, levelsup, _p(out["bow_off"]), _p(out["bow_id"]),
                                              _p(out["bow_val"]), _p(out["fv_off"]), _p(out["fv_node"]), _p(out["fv_feat_off"]),
                                              _p(out["fv_feat"])))
        return out

    def compute_bow_batch(self, desc, counts, levelsup=4):
        """Frame::ComputeBoW (src/Frame.cc:459-466) for a batch: [(BowVector dict, FeatureVector dict)] per image."""
        r = self.compute_bow_batch_raw(desc, counts, levelsup)
        res = []
        for i in range(len(counts)):
            b0, b1 = r["bow_off"][i], r["bow_off"][i + 1]
            bow = {int(k): float(v) for k, v in zip(r["bow_id"][b0:b1], r["bow_val"][b0:b1])}
            fv = {}
            for j in range(r["fv_off"][i], r["fv_off"][i + 1]):
                fv[int(r["fv_node"][j])] = [int(x) for x in r["fv_feat"][r["fv_feat_off"][j]:r["fv_feat_off"][j + 1]]]
            res.append((bow, fv))
        return res

    def transform(self, desc, levelsup=4):
        """transform(features, BowVector, FeatureVector, levelsup) (:1127-1190) for one image."""
        desc = np.ascontiguousarray(desc, dtype=np.uint8).reshape(-1, 32)
        return self.compute_bow_batch(desc, [len(desc)], levelsup)[0]


def stereo_match(left, right, bf, b):
    """Frame::ComputeStereoMatches on the device-resident results of the last single-image extract of both handles."""
    n = left.cap
    uR = np.zeros(n, np.float32)
    depth = np.zeros(n, np.float32)
    nv = C.c_int()
    check(lib().orbb200_stereo_match(left.h, right.h, bf, b, _p(uR), _p(depth), C.byref(nv)))
    return uR, depth, nv.value


# ---- multi-GPU host logic: frames and database rows are sharded per rank --------------------------
def cvt_gray(image, r_first, device=-1):
    """cv::cvtColor(im, gray, CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY) (reference Tracking.cc:177-202)"""
    assert image.dtype == np.uint8 and image.ndim == 3 and image.shape[2] in (3, 4)
    image = np.ascontiguousarray(image)
    h, w, ch = image.shape
    gray = np.zeros((h, w), np.uint8)
    check(lib().orbb200_cvt_gray(_p(image), w, h, image.strides[0], ch, int(bool(r_first)), _p(gray), w, device))
    return gray


def undistort_keypoints(kps, cam, dist_coef, device=-1):
    """Frame::UndistortKeyPoints (reference Frame.cc:470-515): kps [N] KP_DTYPE -> copy with undistorted x, y"""
    kps = np.ascontiguousarray(kps, KP_DTYPE)
    out = np.zeros(len(kps), KP_DTYPE)
    cam = np.ascontiguousarray(cam, np.float32)
    dc = np.ascontiguousarray(dist_coef, np.float32)
    assert cam.size == 4 and dc.size == 5
    check(lib().orbb200_undistort_keypoints(_p(kps), len(kps), _p(cam), _p(dc), _p(out), device))
    return out


class ShardedMatcher:
    """Config 5 on N GPUs (one process per GPU): brute-force top-2 over a row-sharded database with the exchange step
    fused into the search kernel (peer stores over NVLink + flags, include/orbb200.h "config 5 on N GPUs").
    The 64-byte IPC handles are all-gathered once through torch.distributed (host plumbing, any backend)."""
    IPC_HANDLE_BYTES = 64

    def __init__(self, rank, world, nq_max, device=-1, group=None):
        self.L = lib()
        self.h = C.c_void_p()
        self.rank, self.world = rank, world
        handle = np.zeros(self.IPC_HANDLE_BYTES, np.uint8)
        check(self.L.orbb200_exchange_create(C.byref(self.h), rank, world, nq_max, device, _p(handle)))
        if world > 1:
            import torch.distributed as dist
            gathered = [None] * world
            dist.all_gather_object(gathered, handle.tobytes(), group=group)  # also orders every create before any connect
            allh = np.frombuffer(b"".join(gathered), np.uint8).copy()
            check(self.L.orbb200_exchange_connect(self.h, _p(allh)))

    def __del__(self):
        if getattr(self, "h", None) and self.h.value:
            self.L.orbb200_exchange_destroy(self.h)
            self.h = C.c_void_p()

    def match_device(self, q_ptr, nq, db_ptr, ndb_local, index_offset, out_ptr, stream=0):
        """asynchronous; device pointers; every rank calls it the same number of times"""
        check(self.L.orbb200_match_sharded_top2_device(self.h, C.c_void_p(q_ptr), nq, C.c_void_p(db_ptr), ndb_local, index_offset,
                                                        C.c_void_p(out_ptr), C.c_void_p(stream)))

    def status(self, stream=0):
        check(self.L.orbb200_exchange_status(self.h, C.c_void_p(stream)))


def shard_range(n, world, rank):
    """contiguous [lo, hi) slice of n units owned by `rank` of `world` (first n % world ranks get one more)"""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_top2_parts(part, group=None):
    """all-gather of the per-shard top-2 records (config 5's only exchange step): part is a torch int32 tensor
    [nq, 4] (best_dist, best_idx GLOBAL, second_dist, second_idx GLOBAL) on the rank's device (NCCL) or on the CPU
    (gloo); returns [world, nq, 4] ordered by rank = by database position, which the merge's first-seen rule needs."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    out = torch.empty((world * part.shape[0],) + tuple(part.shape[1:]), dtype=part.dtype, device=part.device)
    dist.all_gather_into_tensor(out, part.contiguous(), group=group)  # concatenation along dim 0, rank order
    return out.view((world,) + tuple(part.shape))


class Remap:
    """cv::remap(src, M1, M2, INTER_LINEAR) of the EuRoC example (orbb200_remap_*): maps fixed at construction"""

    def __init__(self, map_x, map_y, device=-1):
        self.mx, self.my = np.ascontiguousarray(map_x, np.float32), np.ascontiguousarray(map_y, np.float32)
        self.dh, self.dw = self.mx.shape
        self.h = C.c_void_p()
        check(lib().orbb200_remap_create(C.byref(self.h), _p(self.mx), _p(self.my), self.dw, self.dh, device))

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orbb200_remap_destroy(self.h)
                self.h = None
        except Exception:  # interpreter shutdown: the module globals may be gone
            pass

    def __call__(self, images):
        """images [batch][h][w] uint8 -> rectified [batch][dh][dw]"""
        images = np.ascontiguousarray(images, np.uint8)
        if images.ndim == 2:
            return self(images[None])[0]
        nb, sh, sw = images.shape
        out = np.zeros((nb, self.dh, self.dw), np.uint8)
        ip_ = (C.c_void_p * nb)(*[images[b].ctypes.data for b in range(nb)])
        op_ = (C.c_void_p * nb)(*[out[b].ctypes.data for b in range(nb)])
        check(lib().orbb200_remap_batch(self.h, ip_, nb, sw, sh, sw, op_, self.dw))
        return out


def init_undistort_rectify_map(K, D, R, P, w, h):
    K = np.ascontiguousarray(np.asarray(K, np.float64)[:3, :3])
    R = np.ascontiguousarray(np.asarray(R, np.float64)[:3, :3])
    P = np.ascontiguousarray(np.asarray(P, np.float64)[:3, :3])
    D = np.ascontiguousarray(np.asarray(D, np.float64).reshape(-1)[:5])
    mx, my = np.zeros((h, w), np.float32), np.zeros((h, w), np.float32)
    check(lib().orbb200_init_undistort_rectify_map(_p(K), _p(D), len(D), _p(R), _p(P), w, h, _p(mx), _p(my)))
    return mx, my
