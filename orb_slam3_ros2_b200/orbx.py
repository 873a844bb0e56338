"""ctypes binding of the C ABI (include/orbx.h) and a host-side mirror of the reference's class surface.

`ORBextractor` mirrors ORB_SLAM3::ORBextractor (ORB_SLAM3/include/ORBextractor.h:43-109): same constructor
arguments, `__call__(image, mask, lapping_area)` returning what operator() produces (monoIndex, keypoints,
descriptors), `mvImagePyramid`, and the Get* accessors.  `DescriptorDistance`, `TH_LOW`, `TH_HIGH` mirror
ORBmatcher (ORB_SLAM3/src/ORBmatcher.cc:35-37, :2058-2074).

There is no CPU fallback: if liborbx.so is missing this module raises at import, and every compute call
fails with OrbxError when no CUDA device is present.  Nothing here imports oracle/.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("ORBX_LIB") or os.path.join(HERE, "liborbx.so")  # ORBX_LIB: a variant build (measurements only)

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

TH_HIGH = 100          # ORBmatcher.cc:35
TH_LOW = 50            # ORBmatcher.cc:36
HISTO_LENGTH = 30      # ORBmatcher.cc:37

ORBX_OK, ORBX_ERR_ARG, ORBX_ERR_CUDA, ORBX_ERR_CAPACITY, ORBX_ERR_NCCL, ORBX_ERR_STATE = 0, -1, -2, -3, -4, -5


class OrbxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"orbx error {code}: {msg}")
        self.code = code


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int), ("ini_th_fast", C.c_int),
                ("min_th_fast", C.c_int), ("device", C.c_int), ("fp_mode", C.c_int)]


_u8p = C.POINTER(C.c_uint8)
_i32p = C.POINTER(C.c_int32)
_i64p = C.POINTER(C.c_int64)
_f32p = C.POINTER(C.c_float)
_vp = C.c_void_p

# every symbol include/orbx.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "orbx_last_error": (C.c_char_p, []),
    "orbx_version": (C.c_int, []),
    "orbx_launch_count": (C.c_ulonglong, []),
    "orbx_create": (C.c_int, [C.POINTER(Params), C.POINTER(_vp)]),
    "orbx_destroy": (None, [_vp]),
    "orbx_get_tables": (C.c_int, [_vp, _f32p, _f32p, _f32p, _f32p, _i32p]),
    "orbx_extract": (C.c_int, [_vp, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _u8p, C.c_int, _i32p, _i32p]),
    "orbx_extract_color": (C.c_int, [_vp, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp,
                                     _u8p, C.c_int, _i32p, _i32p]),
    "orbx_get_staged_gray": (C.c_int, [_vp, _u8p, _i32p, _i32p]),
    "orbx_get_pyramid": (C.c_int, [_vp, C.c_int, C.c_int, _u8p, _i32p, _i32p]),
    "orbx_extract_batch": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, _vp, _vp,
                                     C.c_int, _vp, _vp]),
    "orbx_extract_batch_async": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, _vp, _vp,
                                           C.c_int, _vp, _vp]),
    "orbx_check": (C.c_int, [_vp]),
    "orbx_extract_batch_device": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int]),
    "orbx_batch_results": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), _i32p]),
    "orbx_batch_download": (C.c_int, [_vp, C.c_int, _vp, _u8p, C.c_int, _i32p, _i32p]),
    "orbx_batch_counts": (C.c_int, [_vp, _i32p, _i32p]),
    "orbx_sync": (C.c_int, [_vp]),
    "orbx_set_subbatches": (C.c_int, [_vp, C.c_int]),
    "orbx_get_stream": (_vp, [_vp]),
    "orbx_last_timing": (C.c_int, [_vp, _f32p, _f32p]),
    "orbx_get_candidates": (C.c_int, [_vp, C.c_int, C.c_int, _i32p, C.c_int, _i32p]),
    "orbx_get_blurred": (C.c_int, [_vp, C.c_int, C.c_int, _u8p, _i32p, _i32p]),
    "orbx_synth_frames_device": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_int]),
    "orbx_hamming": (C.c_int, [_u8p, _u8p]),
    "orbx_match_top2": (C.c_int, [_u8p, C.c_int, _u8p, C.c_int64, _i64p, _i32p, _i64p, _i32p, C.c_int]),
    "orbx_match_top2_device": (C.c_int, [_vp, C.c_int, _vp, C.c_int64, C.c_int64, _vp, _vp, _vp, _vp, _vp]),
    "orbx_match_top2_device_noalloc": (C.c_int, [_vp, C.c_int, _vp, C.c_int64, C.c_int64, _vp, _vp, _vp, _vp, _vp]),
    "orbx_match_batch_consecutive": (C.c_int, [_vp, _vp, _vp, _vp, _vp, C.c_int]),
    "orbx_match_stereo_banded": (C.c_int, [_vp, _u8p, C.c_int, _vp, _u8p, C.c_int, _f32p, C.c_int, C.c_int, C.c_float,
                                           _i32p, _i32p, C.c_int]),
    "orbx_accept_top2": (C.c_int, [C.c_int, C.c_float, _i32p, _i32p, _i64p, _i32p, _i32p, C.c_int, _u8p, C.c_int]),
    "orbx_accept_top2_device": (C.c_int, [C.c_int, C.c_float, _vp, _vp, _vp, _vp, _vp, C.c_int, _vp, _vp]),
    "orbx_search_grid_top2": (C.c_int, [_vp, _u8p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, _u8p, _f32p, _vp, _u8p,
                                        C.c_int, _i32p, _i32p, _i32p, _i32p, _i32p, _i32p, C.c_int]),
    "orbx_compute_stereo_matches": (C.c_int, [_vp, _vp, C.c_int, _vp, _u8p, C.c_int, _vp, _u8p, C.c_int, C.c_float, C.c_float,
                                              _f32p, _f32p]),
    "orbx_compute_stereo_matches_batch": (C.c_int, [_vp, _vp, C.c_float, C.c_float, _f32p, _f32p]),
    "orbx_undistort_keypoints": (C.c_int, [_vp, C.c_int, _f32p, _f32p, C.c_int, _f32p, _vp, C.c_int]),
    "orbx_image_bounds": (C.c_int, [C.c_int, C.c_int, _f32p, _f32p, C.c_int, _f32p, _f32p, C.c_int]),
    "orbx_undistort_batch": (C.c_int, [_vp, _f32p, _f32p, C.c_int, _f32p, _vp, C.POINTER(_vp)]),
    "orbx_search_for_initialization": (C.c_int, [_vp, _u8p, C.c_int, _vp, _u8p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                                 _f32p, C.c_int, C.c_float, C.c_int, _i32p, _i32p, C.c_int]),
    "orbx_search_for_initialization_batch": (C.c_int, [_vp, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_float,
                                                       C.c_int, _i32p, _i32p, _f32p]),
    "orbx_search_by_bow": (C.c_int, [_u8p, C.c_int, _u8p, _f32p, _i32p, _i32p, _i32p, C.c_int, _u8p, C.c_int, _f32p, _i32p, _i32p, _i32p,
                                     C.c_int, C.c_float, C.c_int, _i32p, _i32p, C.c_int]),
    "orbx_search_by_bow_kf": (C.c_int, [_u8p, C.c_int, _u8p, _f32p, _i32p, _i32p, _i32p, C.c_int, _u8p, C.c_int, _u8p, _f32p, _i32p, _i32p,
                                        _i32p, C.c_int, C.c_float, C.c_int, _i32p, _i32p, C.c_int]),
    "orbx_capacity": (C.c_int, [_vp, C.c_int, C.c_int, _i32p]),
    "orbx_thread_cleanup": (None, []),
    "orbx_search_by_projection_map": (C.c_int, [_vp, _i32p, _u8p, _u8p, _i32p, _vp, _u8p, _vp, _vp, _i32p, C.c_int, C.c_int, C.c_int, _vp,
                                                C.c_float, C.c_int, C.c_float, _i32p, _i32p, C.c_int]),
    "orbx_search_by_projection": (C.c_int, [_vp, _u8p, _vp, _u8p, _i32p, _vp, _u8p, _vp, _vp, _i32p, _vp, C.c_int, C.c_int, C.c_int, _vp,
                                            _i32p, _i32p, C.c_int]),
    "orbx_debug_check_guards": (C.c_int, [_vp]),
    "orbx_extract_frame": (C.c_int, [_vp, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp]),
    "orbx_bench_popc": (C.c_int, [C.c_int, C.POINTER(C.c_double)]),
    "orbx_voc_create": (C.c_int, [C.c_int, _i32p, _u8p, C.POINTER(C.c_double), C.c_int, C.c_int, C.POINTER(_vp)]),
    "orbx_voc_destroy": (None, [_vp]),
    "orbx_voc_info": (C.c_int, [_vp, _i32p, _i32p, _i32p, _i32p]),
    "orbx_bow_transform": (C.c_int, [_vp, _u8p, _i32p, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "orbx_bow_transform_extracted": (C.c_int, [_vp, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "orbx_comm_unique_id": (C.c_int, [_u8p]),
    "orbx_comm_init": (C.c_int, [C.c_int, C.c_int, _u8p, C.c_int, C.POINTER(_vp)]),
    "orbx_comm_destroy": (None, [_vp]),
    "orbx_match_top2_sharded": (C.c_int, [_vp, _vp, C.c_int, _vp, C.c_int64, C.c_int64, _vp, _vp, _vp, _vp]),
    "orbx_match_top2_sharded_async": (C.c_int, [_vp, _vp, C.c_int, _vp, C.c_int64, C.c_int64, _vp, _vp, _vp, _vp]),
    "orbx_comm_sync": (C.c_int, [_vp]),
    "orbx_merge_top2_device": (C.c_int, [_vp, _vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp]),
}

_lib = None


def lib():
    """The loaded liborbx.so.  Raises (no fallback) if the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: build it with `make -C orb_slam3_ros2_b200/csrc` "
                              "(or __graft_entry__.build()); this package has no CPU path")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)  # AttributeError = header/library mismatch
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def _check(rc):
    if rc != ORBX_OK:
        raise OrbxError(rc, lib().orbx_last_error().decode(errors="replace"))


def _ptr(a, t=_u8p):
    return a.ctypes.data_as(t)


def launch_count():
    return int(lib().orbx_launch_count())


class ORBextractor:
    """Mirror of ORB_SLAM3::ORBextractor (ORBextractor.h:43-109) over the C ABI."""
    HARRIS_SCORE, FAST_SCORE = 0, 1  # ORBextractor.h:47

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, device=-1, fp_mode="strict"):
        self._h = _vp()
        self.nfeatures, self.nlevels = int(nfeatures), int(nlevels)
        p = Params(int(nfeatures), float(scaleFactor), int(nlevels), int(iniThFAST), int(minThFAST), int(device),
                   0 if fp_mode == "strict" else 1)
        _check(lib().orbx_create(C.byref(p), C.byref(self._h)))
        f = [np.zeros(self.nlevels, dtype=np.float32) for _ in range(4)]
        self._nfeat = np.zeros(self.nlevels, dtype=np.int32)
        _check(lib().orbx_get_tables(self._h, *[_ptr(a, _f32p) for a in f], _ptr(self._nfeat, _i32p)))
        self._sf, self._isf, self._sig2, self._isig2 = f
        self._scale_factor = np.float32(np.float64(np.float32(scaleFactor)))
        self.mvImagePyramid = []
        self._w = self._h_img = 0
        self._frames = 0

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().orbx_destroy(self._h)
            self._h = _vp()

    __del__ = close

    # --- accessors (ORBextractor.h:61-83)
    def GetLevels(self): return self.nlevels
    def GetScaleFactor(self): return float(self._scale_factor)
    def GetScaleFactors(self): return self._sf.copy()
    def GetInverseScaleFactors(self): return self._isf.copy()
    def GetScaleSigmaSquares(self): return self._sig2.copy()
    def GetInverseScaleSigmaSquares(self): return self._isig2.copy()
    def features_per_level(self): return self._nfeat.copy()

    @property
    def capacity(self):
        """Upper bound of keypoints per frame for the size of the last frames seen (orbx_capacity: sum over the levels of
        max(N_l + 4, 4 * nIni_l + 1)); before any frame, the size-independent part sum(N_l + 4)."""
        return self._cap if getattr(self, "_cap", 0) else int(self._nfeat.sum()) + 4 * self.nlevels

    def capacity_for(self, w, h):
        c = C.c_int32(0)
        _check(lib().orbx_capacity(self._h, int(w), int(h), C.byref(c)))
        self._cap = int(c.value)
        return self._cap

    # --- operator() (ORBextractor.cc:1086-1168)
    def __call__(self, image, mask=None, vLappingArea=(0, 1000), pyramid=True):
        """-> (monoIndex, keypoints[KP_DTYPE], descriptors[n,32]).  `mask` is ignored, as in the reference."""
        cap = self.capacity if image is None or image.size == 0 else self.capacity_for(image.shape[1], image.shape[0])
        kps = np.zeros(cap, dtype=KP_DTYPE)
        desc = np.zeros((cap, 32), dtype=np.uint8)
        n, mono = C.c_int32(0), C.c_int32(0)
        if image is None or image.size == 0:
            _check(lib().orbx_extract(self._h, None, 0, 0, 0, vLappingArea[0], vLappingArea[1], None, None, 0,
                                      C.byref(n), C.byref(mono)))
            return mono.value, kps[:0], desc[:0]
        assert image.dtype == np.uint8 and image.ndim == 2  # assert(image.type() == CV_8UC1), ORBextractor.cc:1094
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        h, w = image.shape
        _check(lib().orbx_extract(self._h, _ptr(image), w, h, image.strides[0], vLappingArea[0], vLappingArea[1],
                                  kps.ctypes.data_as(_vp), _ptr(desc), cap, C.byref(n), C.byref(mono)))
        self._frames = 1
        if pyramid:
            self.mvImagePyramid = [self.pyramid_level(0, l)[19:-19, 19:-19] for l in range(self.nlevels)]
        return mono.value, kps[:n.value].copy(), desc[:n.value].copy()

    def check_guards(self):
        """ORBX_GUARD=1 only: raises if a kernel wrote behind one of this handle's device buffers."""
        _check(lib().orbx_debug_check_guards(self._h))

    def set_subbatches(self, n):
        _check(lib().orbx_set_subbatches(self._h, int(n)))

    def extract_color(self, image, rgb=True, new_size=None, vLappingArea=(0, 1000)):
        """System::Track*'s resize (System.cc:420-425) + Tracking::GrabImage*'s cvtColor (Tracking.cc:1569-1582) fused
        ahead of operator().  image: (H, W), (H, W, 3) or (H, W, 4) uint8; new_size = (w, h) or None.
        -> (monoIndex, keypoints, descriptors); the grey frame the extractor saw is `staged_gray()`."""
        image = np.ascontiguousarray(image, dtype=np.uint8)
        ch = 1 if image.ndim == 2 else image.shape[2]
        h, w = image.shape[:2]
        nw, nh = new_size if new_size else (0, 0)
        cap = self.capacity_for(nw or w, nh or h)
        kps = np.zeros(cap, dtype=KP_DTYPE)
        desc = np.zeros((cap, 32), dtype=np.uint8)
        n, mono = C.c_int32(0), C.c_int32(0)
        _check(lib().orbx_extract_color(self._h, _ptr(image), w, h, image.strides[0], ch, 1 if rgb else 0, int(nw), int(nh),
                                        vLappingArea[0], vLappingArea[1], kps.ctypes.data_as(_vp), _ptr(desc), cap, C.byref(n),
                                        C.byref(mono)))
        self._frames = 1
        return mono.value, kps[:n.value].copy(), desc[:n.value].copy()

    def staged_gray(self):
        w, h = C.c_int32(0), C.c_int32(0)
        _check(lib().orbx_get_staged_gray(self._h, None, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), dtype=np.uint8)
        _check(lib().orbx_get_staged_gray(self._h, _ptr(out), None, None))
        return out

    def pyramid_level(self, frame, level):
        """Padded plane (h+38, w+38) of `level`; mvImagePyramid[level] is its [19:-19, 19:-19] view."""
        w, h = C.c_int32(0), C.c_int32(0)
        _check(lib().orbx_get_pyramid(self._h, frame, level, None, C.byref(w), C.byref(h)))
        out = np.zeros((h.value + 38, w.value + 38), dtype=np.uint8)
        _check(lib().orbx_get_pyramid(self._h, frame, level, _ptr(out), None, None))
        return out

    def blurred_level(self, frame, level):
        w, h = C.c_int32(0), C.c_int32(0)
        _check(lib().orbx_get_blurred(self._h, frame, level, None, C.byref(w), C.byref(h)))
        out = np.zeros((h.value, w.value), dtype=np.uint8)
        _check(lib().orbx_get_blurred(self._h, frame, level, _ptr(out), None, None))
        return out

    def candidates(self, frame, level):
        n = C.c_int32(0)
        _check(lib().orbx_get_candidates(self._h, frame, level, None, 0, C.byref(n)))
        out = np.zeros((max(n.value, 1), 3), dtype=np.int32)
        _check(lib().orbx_get_candidates(self._h, frame, level, _ptr(out, _i32p), n.value, C.byref(n)))
        return out[:n.value]

    # --- batch path (frames are independent: ORBextractor holds only constants + scratch)
    def extract_batch(self, images, vLappingArea=(0, 1000), download=True):
        """images: (F, H, W) uint8 host array.  -> (mono[F], n[F], kps[F,cap], desc[F,cap,32])"""
        images = np.ascontiguousarray(images, dtype=np.uint8)
        F, h, w = images.shape
        cap = self.capacity_for(w, h)
        n = np.zeros(F, dtype=np.int32)
        mono = np.zeros(F, dtype=np.int32)
        kps = np.zeros((F, cap), dtype=KP_DTYPE) if download else None
        desc = np.zeros((F, cap, 32), dtype=np.uint8) if download else None
        _check(lib().orbx_extract_batch(self._h, images.ctypes.data_as(_vp), F, w, h, w, w * h, vLappingArea[0],
                                        vLappingArea[1], kps.ctypes.data_as(_vp) if download else None,
                                        desc.ctypes.data_as(_vp) if download else None, cap, n.ctypes.data_as(_vp),
                                        mono.ctypes.data_as(_vp)))
        self._frames = F
        return mono, n, kps, desc

    def extract_batch_ptr(self, host_ptr, F, w, h, vLappingArea, kps_ptr, desc_ptr, cap, n_ptr, mono_ptr):
        """Raw-pointer form (pinned host buffers owned by the caller; used by bench.py's e2e leg)."""
        _check(lib().orbx_extract_batch(self._h, host_ptr, F, w, h, w, w * h, vLappingArea[0], vLappingArea[1], kps_ptr,
                                        desc_ptr, cap, n_ptr, mono_ptr))
        self._frames = F

    def extract_batch_async_ptr(self, host_ptr, F, w, h, vLappingArea, kps_ptr, desc_ptr, cap, n_ptr, mono_ptr):
        """Enqueue-only form (pinned host buffers): H2D + kernels + D2H on this handle's stream; call check() later."""
        _check(lib().orbx_extract_batch_async(self._h, host_ptr, F, w, h, w, w * h, vLappingArea[0], vLappingArea[1],
                                              kps_ptr, desc_ptr, cap, n_ptr, mono_ptr))
        self._frames = F

    def match_consecutive_async_ptr(self, idx1, d1, idx2, d2):
        _check(lib().orbx_match_batch_consecutive(self._h, idx1, d1, idx2, d2, 2))

    def check(self):
        """Wait for the handle's stream and raise if a device-side capacity invariant tripped."""
        _check(lib().orbx_check(self._h))

    def extract_batch_device(self, dev_ptr, F, w, h, vLappingArea=(0, 1000), stride=None, frame_stride=None):
        _check(lib().orbx_extract_batch_device(self._h, dev_ptr, F, w, h, stride or w, frame_stride or w * h,
                                               vLappingArea[0], vLappingArea[1]))
        self._frames = F
        self.capacity_for(w, h)

    def sync(self):
        _check(lib().orbx_sync(self._h))

    def stream_ptr(self):
        return lib().orbx_get_stream(self._h)

    def timing(self):
        """-> (total_ms, [pyramid, fast, quadtree, blur, describe] ms) of the last batch (CUDA events)."""
        tot = C.c_float(0)
        st = (C.c_float * 5)()
        _check(lib().orbx_last_timing(self._h, C.byref(tot), st))
        return tot.value, [st[i] for i in range(5)]

    def download(self, frame):
        cap = self.capacity
        kps = np.zeros(cap, dtype=KP_DTYPE)
        desc = np.zeros((cap, 32), dtype=np.uint8)
        n, mono = C.c_int32(0), C.c_int32(0)
        _check(lib().orbx_batch_download(self._h, frame, kps.ctypes.data_as(_vp), _ptr(desc), cap, C.byref(n),
                                         C.byref(mono)))
        return mono.value, kps[:n.value].copy(), desc[:n.value].copy()

    def counts(self):
        """-> (n, mono) int32 arrays over all frames of the last batch."""
        n, mono = np.zeros(self._frames, np.int32), np.zeros(self._frames, np.int32)
        _check(lib().orbx_batch_counts(self._h, _ptr(n, _i32p), _ptr(mono, _i32p)))
        return n, mono

    def results_device(self):
        """-> (d_kps, d_desc, d_n, d_mono, cap) raw device addresses of the last batch."""
        p = [_vp() for _ in range(4)]
        cap = C.c_int32(0)
        _check(lib().orbx_batch_results(self._h, *[C.byref(x) for x in p], C.byref(cap)))
        return p[0].value, p[1].value, p[2].value, p[3].value, cap.value

    def match_consecutive(self):
        """Top-2 of frame f against frame f+1 over the last batch -> idx1, d1, idx2, d2, each (F-1, cap) int32."""
        F = self._frames
        cap = self.results_device()[4]
        out = [np.zeros((max(F - 1, 0), cap), dtype=np.int32) for _ in range(4)]
        if F >= 2:
            _check(lib().orbx_match_batch_consecutive(self._h, *[o.ctypes.data_as(_vp) for o in out], 1))
        return out

    def match_consecutive_device(self, d_idx1, d_d1, d_idx2, d_d2):
        _check(lib().orbx_match_batch_consecutive(self._h, d_idx1, d_d1, d_idx2, d_d2, 0))


def DescriptorDistance(a, b):
    """ORBmatcher::DescriptorDistance (ORBmatcher.cc:2058-2074)."""
    a = np.ascontiguousarray(a, dtype=np.uint8)
    b = np.ascontiguousarray(b, dtype=np.uint8)
    assert a.size == 32 and b.size == 32
    return lib().orbx_hamming(_ptr(a), _ptr(b))


def match_top2(q, db, device=-1):
    """Brute-force k=2 (knnMatch order) -> idx1 (int64), d1, idx2, d2."""
    q = np.ascontiguousarray(q, dtype=np.uint8).reshape(-1, 32)
    db = np.ascontiguousarray(db, dtype=np.uint8).reshape(-1, 32)
    nq = len(q)
    i1, i2 = np.zeros(nq, dtype=np.int64), np.zeros(nq, dtype=np.int64)
    d1, d2 = np.zeros(nq, dtype=np.int32), np.zeros(nq, dtype=np.int32)
    _check(lib().orbx_match_top2(_ptr(q), nq, _ptr(db), len(db), _ptr(i1, _i64p), _ptr(d1, _i32p), _ptr(i2, _i64p),
                                 _ptr(d2, _i32p), device))
    return i1, d1, i2, d2


def match_stereo_banded(kps_l, desc_l, kps_r, desc_r, scale_factors, n_rows, max_d, device=-1):
    """Hamming stage of Frame::ComputeStereoMatches (Frame.cc:811-894) -> best_idx, best_dist."""
    kps_l = np.ascontiguousarray(kps_l, dtype=KP_DTYPE)
    kps_r = np.ascontiguousarray(kps_r, dtype=KP_DTYPE)
    desc_l = np.ascontiguousarray(desc_l, dtype=np.uint8)
    desc_r = np.ascontiguousarray(desc_r, dtype=np.uint8)
    sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
    bi = np.zeros(len(kps_l), dtype=np.int32)
    bd = np.zeros(len(kps_l), dtype=np.int32)
    _check(lib().orbx_match_stereo_banded(kps_l.ctypes.data_as(_vp), _ptr(desc_l), len(kps_l), kps_r.ctypes.data_as(_vp),
                                          _ptr(desc_r), len(kps_r), _ptr(sf, _f32p), len(sf), int(n_rows), float(max_d),
                                          _ptr(bi, _i32p), _ptr(bd, _i32p), device))
    return bi, bd


def bench_popc(device=-1):
    """Measured sustained POPC rate (giga-POPC/s) of the device: the matcher's roofline denominator."""
    r = C.c_double(0)
    _check(lib().orbx_bench_popc(device, C.byref(r)))
    return r.value


# orbx_accept_rule (include/orbx.h)
ACCEPT_PROJ_MAP, ACCEPT_BOW_KF_F, ACCEPT_INIT, ACCEPT_BOW_KF_KF, ACCEPT_BEST_HIGH, ACCEPT_BEST_LOW, ACCEPT_KNN_RATIO, ACCEPT_LOW_RATIO = range(8)

GRID_QUERY_DTYPE = np.dtype([("u", "<f4"), ("v", "<f4"), ("r", "<f4"), ("min_level", "<i4"), ("max_level", "<i4"),
                             ("ur", "<f4")])
assert GRID_QUERY_DTYPE.itemsize == 24


def _opt(a, dtype, t):
    if a is None:
        return None, None
    a = np.ascontiguousarray(a, dtype=dtype)
    return a, _ptr(a, t)


def accept_top2(rule, d1, d2, nnratio=0.7, idx2=None, level1=None, level2=None, device=-1):
    """The reference's accept rules on top-2 results, evaluated on the device (ORBmatcher.cc:123-128, :327-329, :697-699,
    :848-850, :1767, Frame.cc:1149) -> bool array."""
    d1 = np.ascontiguousarray(d1, dtype=np.int32)
    d2 = np.ascontiguousarray(d2, dtype=np.int32)
    n = len(d1)
    k2, p2 = _opt(idx2, np.int64, _i64p)
    l1, pl1 = _opt(level1, np.int32, _i32p)
    l2, pl2 = _opt(level2, np.int32, _i32p)
    acc = np.zeros(max(n, 1), dtype=np.uint8)
    _check(lib().orbx_accept_top2(int(rule), float(nnratio), _ptr(d1, _i32p), _ptr(d2, _i32p), p2, pl1, pl2, n, _ptr(acc), device))
    return acc[:n].astype(bool)


def search_grid_top2(kps_t, desc_t, bounds, queries, desc_q, skip_t=None, u_right_t=None, device=-1):
    """Guided search: Frame::GetFeaturesInArea windows (Frame.cc:657-723) + the strict-< best/second loop
    (ORBmatcher.cc:75-117) -> idx1, d1, idx2, d2, lvl1, lvl2 (int32).  bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    kps_t = np.ascontiguousarray(kps_t, dtype=KP_DTYPE)
    desc_t = np.ascontiguousarray(desc_t, dtype=np.uint8).reshape(-1, 32)
    queries = np.ascontiguousarray(queries, dtype=GRID_QUERY_DTYPE)
    desc_q = np.ascontiguousarray(desc_q, dtype=np.uint8).reshape(-1, 32)
    nq = len(queries)
    sk, psk = _opt(skip_t, np.uint8, _u8p)
    ur, pur = _opt(u_right_t, np.float32, _f32p)
    out = [np.zeros(max(nq, 1), dtype=np.int32) for _ in range(6)]
    _check(lib().orbx_search_grid_top2(kps_t.ctypes.data_as(_vp), _ptr(desc_t), len(kps_t), *[float(b) for b in bounds], psk, pur,
                                       queries.ctypes.data_as(_vp), _ptr(desc_q), nq, *[_ptr(o, _i32p) for o in out], device))
    return [o[:nq] for o in out]


def compute_stereo_matches(left, right, kps_l, desc_l, kps_r, desc_r, bf, b, frame=0):
    """Frame::ComputeStereoMatches (Frame.cc:811-980) on the pyramids held by two ORBextractor objects whose last
    extraction produced kps_l / kps_r -> (mvuRight, mvDepth)."""
    kps_l = np.ascontiguousarray(kps_l, dtype=KP_DTYPE)
    kps_r = np.ascontiguousarray(kps_r, dtype=KP_DTYPE)
    desc_l = np.ascontiguousarray(desc_l, dtype=np.uint8)
    desc_r = np.ascontiguousarray(desc_r, dtype=np.uint8)
    n = len(kps_l)
    ur = np.zeros(max(n, 1), dtype=np.float32)
    dp = np.zeros(max(n, 1), dtype=np.float32)
    _check(lib().orbx_compute_stereo_matches(left._h, right._h, frame, kps_l.ctypes.data_as(_vp), _ptr(desc_l), n,
                                             kps_r.ctypes.data_as(_vp), _ptr(desc_r), len(kps_r), float(bf), float(b),
                                             _ptr(ur, _f32p), _ptr(dp, _f32p)))
    return ur[:n], dp[:n]


def compute_stereo_matches_batch(left, right, bf, b, out=None):
    """Frame::ComputeStereoMatches for every pair of the two extractors' last batches, all inputs read on the device ->
    (mvuRight, mvDepth), each (F, cap) float32 (entries past a frame's keypoint count are unspecified)."""
    F, cap = left._frames, left.results_device()[4]
    ur, dp = out if out is not None else (np.zeros((F, cap), np.float32), np.zeros((F, cap), np.float32))
    _check(lib().orbx_compute_stereo_matches_batch(left._h, right._h, float(bf), float(b), _ptr(ur, _f32p), _ptr(dp, _f32p)))
    return ur, dp


def _camera(K, dist, P):
    K = np.ascontiguousarray(K, dtype=np.float32).reshape(3, 3)
    P = K if P is None else np.ascontiguousarray(P, dtype=np.float32).reshape(3, 3)
    return K, np.ascontiguousarray(dist, dtype=np.float32).ravel(), P


def undistort_keypoints(kps, K, dist, P=None, device=-1):
    """Frame::UndistortKeyPoints (Frame.cc:747-780) for a pinhole camera: mvKeys -> mvKeysUn.  K = Pinhole::toK(), P = mK
    (default K), dist = mDistCoef."""
    K, dist, P = _camera(K, dist, P)
    kps = np.ascontiguousarray(kps, dtype=KP_DTYPE)
    out = np.zeros(max(len(kps), 1), dtype=KP_DTYPE)
    _check(lib().orbx_undistort_keypoints(kps.ctypes.data_as(_vp), len(kps), _ptr(K, _f32p), _ptr(dist, _f32p), len(dist),
                                          _ptr(P, _f32p), out.ctypes.data_as(_vp), device))
    return out[:len(kps)]


def image_bounds(w, h, K, dist, P=None, device=-1):
    """Frame::ComputeImageBounds (Frame.cc:782-809) -> (mnMinX, mnMaxX, mnMinY, mnMaxY) float32."""
    K, dist, P = _camera(K, dist, P)
    out = np.zeros(4, np.float32)
    _check(lib().orbx_image_bounds(int(w), int(h), _ptr(K, _f32p), _ptr(dist, _f32p), len(dist), _ptr(P, _f32p), _ptr(out, _f32p),
                                   device))
    return out


def undistort_batch(extractor, K, dist, P=None, download=True):
    """mvKeysUn of every frame of the extractor's last batch, computed in place on the device -> (F, cap) KP_DTYPE array
    (entries past a frame's keypoint count are unspecified) or, with download=False, the device pointer."""
    K, dist, P = _camera(K, dist, P)
    F, cap = extractor._frames, extractor.capacity
    out = np.zeros((F, cap), dtype=KP_DTYPE) if download else None
    dptr = _vp()
    _check(lib().orbx_undistort_batch(extractor._h, _ptr(K, _f32p), _ptr(dist, _f32p), len(dist), _ptr(P, _f32p),
                                      out.ctypes.data_as(_vp) if download else None, C.byref(dptr)))
    return out if download else dptr.value


def search_for_initialization(kps1, desc1, kps2, desc2, bounds, prev_matched, window_size=100, nnratio=0.9, check_orientation=True,
                              device=-1):
    """ORBmatcher::SearchForInitialization (ORBmatcher.cc:648-763) -> (nmatches, vnMatches12, vbPrevMatched after the call).
    kps1 / kps2 = mvKeysUn of the two frames, bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY) of the second frame's grid."""
    kps1 = np.ascontiguousarray(kps1, dtype=KP_DTYPE)
    kps2 = np.ascontiguousarray(kps2, dtype=KP_DTYPE)
    desc1 = np.ascontiguousarray(desc1, dtype=np.uint8).reshape(-1, 32)
    desc2 = np.ascontiguousarray(desc2, dtype=np.uint8).reshape(-1, 32)
    prev = np.array(prev_matched, dtype=np.float32).reshape(-1, 2).copy()
    if len(prev) != len(kps1):
        raise ValueError("prev_matched must hold one point per keypoint of the first frame")
    m12 = np.zeros(max(len(kps1), 1), dtype=np.int32)
    n = np.zeros(1, dtype=np.int32)
    _check(lib().orbx_search_for_initialization(kps1.ctypes.data_as(_vp), _ptr(desc1), len(kps1), kps2.ctypes.data_as(_vp), _ptr(desc2),
                                                len(kps2), *[float(b) for b in bounds], _ptr(prev, _f32p), int(window_size),
                                                float(nnratio), int(bool(check_orientation)), _ptr(m12, _i32p), _ptr(n, _i32p), device))
    return int(n[0]), m12[:len(kps1)], prev


def search_for_initialization_batch(extractor, bounds, window_size=100, nnratio=0.9, check_orientation=True, use_undistorted=False):
    """SearchForInitialization of every consecutive pair (frame p -> p + 1) of the extractor's last batch, read on the device ->
    (nmatches[F-1], vnMatches12[F-1, cap], vbPrevMatched[F-1, cap, 2])."""
    F, cap = extractor._frames, extractor.capacity
    m12 = np.zeros((max(F - 1, 1), cap), dtype=np.int32)
    n = np.zeros(max(F - 1, 1), dtype=np.int32)
    prev = np.zeros((max(F - 1, 1), cap, 2), dtype=np.float32)
    _check(lib().orbx_search_for_initialization_batch(extractor._h, int(bool(use_undistorted)), *[float(b) for b in bounds],
                                                      int(window_size), float(nnratio), int(bool(check_orientation)),
                                                      _ptr(m12, _i32p), _ptr(n, _i32p), _ptr(prev, _f32p)))
    return n[:F - 1], m12[:F - 1], prev[:F - 1]


class FrameLevel(C.Structure):
    _fields_ = [("w", C.c_int), ("h", C.c_int), ("pitch", C.c_int), ("offset", C.c_uint)]


class FrameResult(C.Structure):
    """orbx_frame_result (include/orbx.h)."""
    _fields_ = [("kps", _vp), ("desc", _vp), ("n", C.c_int), ("mono_index", C.c_int), ("pyramid", _vp), ("nlevels", C.c_int),
                ("level", FrameLevel * 16)]


def extract_frame(extractor, image, vLappingArea=(0, 1000), pyramid=True):
    """orbx_extract_frame: operator() with one host synchronisation (the call compat/ORBextractor.cc makes).
    -> (monoIndex, keypoints, descriptors, [padded planes] or None); the arrays are copies of the handle's staging buffer."""
    r = FrameResult()
    if image is None or image.size == 0:
        _check(lib().orbx_extract_frame(extractor._h, None, 0, 0, 0, vLappingArea[0], vLappingArea[1], 1 if pyramid else 0, C.byref(r)))
        return r.mono_index, np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8), None
    assert image.dtype == np.uint8 and image.ndim == 2
    if image.strides[1] != 1:
        image = np.ascontiguousarray(image)
    h, w = image.shape
    _check(lib().orbx_extract_frame(extractor._h, C.cast(image.ctypes.data, _u8p), w, h, image.strides[0], vLappingArea[0], vLappingArea[1],
                                    1 if pyramid else 0, C.byref(r)))
    extractor._frames = 1
    n = r.n
    kps = np.frombuffer((C.c_char * (n * 28)).from_address(r.kps), dtype=KP_DTYPE).copy() if n else np.zeros(0, KP_DTYPE)
    desc = np.frombuffer((C.c_char * (n * 32)).from_address(r.desc), dtype=np.uint8).reshape(n, 32).copy() if n else np.zeros((0, 32), np.uint8)
    planes = None
    if pyramid and r.pyramid:
        planes = []
        for l in range(r.nlevels):
            L = r.level[l]
            raw = np.frombuffer((C.c_char * (L.pitch * (L.h + 38))).from_address(r.pyramid + L.offset), dtype=np.uint8)
            planes.append(raw.reshape(L.h + 38, L.pitch)[:, :L.w + 38].copy())
    return r.mono_index, kps, desc, planes


class ProjParams(C.Structure):
    """orbx_proj_params (include/orbx.h)."""
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float), ("min_x", C.c_float), ("max_x", C.c_float),
                ("min_y", C.c_float), ("max_y", C.c_float), ("scale_factors", C.c_float * 16), ("nlevels", C.c_int), ("mbf", C.c_float),
                ("mb", C.c_float), ("th", C.c_float), ("mono", C.c_int), ("check_orientation", C.c_int)]


def search_by_projection(cases, cam, bounds, scale_factors, th, mono, mbf=0.0, mb=0.0, check_orientation=True, device=-1):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (ORBmatcher.cc:1676-1887) on the device for a list of
    independent frame pairs.  Each case is a dict with kps_last, last_flags, world, mp_desc, kps_cur, desc_cur, pose_cw, pose_lw and
    optionally cur_u_right, cur_taken (None on every case or on none).  -> [(nmatches, match_cur[n_cur])] per case."""
    P = len(cases)
    cap_l = max([len(c["kps_last"]) for c in cases] + [1])
    cap_c = max([len(c["kps_cur"]) for c in cases] + [1])
    kl = np.zeros((P, cap_l), dtype=KP_DTYPE); fl = np.zeros((P, cap_l), np.uint8); wd = np.zeros((P, cap_l, 3), np.float32)
    md = np.zeros((P, cap_l, 32), np.uint8); kc = np.zeros((P, cap_c), dtype=KP_DTYPE); dc = np.zeros((P, cap_c, 32), np.uint8)
    has_ur = cases[0].get("cur_u_right") is not None
    has_tk = cases[0].get("cur_taken") is not None
    ur = np.full((P, cap_c), -1, np.float32); tk = np.zeros((P, cap_c), np.uint8)
    nl = np.zeros(P, np.int32); nc = np.zeros(P, np.int32); poses = np.zeros((P, 14), np.float32)
    for p, c in enumerate(cases):
        a, b = len(c["kps_last"]), len(c["kps_cur"])
        nl[p], nc[p] = a, b
        kl[p, :a] = c["kps_last"]; fl[p, :a] = c["last_flags"]; wd[p, :a] = np.asarray(c["world"], np.float32).reshape(-1, 3)
        md[p, :a] = c["mp_desc"]; kc[p, :b] = c["kps_cur"]; dc[p, :b] = c["desc_cur"]
        if has_ur: ur[p, :b] = c["cur_u_right"]
        if has_tk: tk[p, :b] = c["cur_taken"]
        poses[p, :7] = c["pose_cw"]; poses[p, 7:] = c["pose_lw"]
    prm = ProjParams()
    prm.fx, prm.fy, prm.cx, prm.cy = [float(v) for v in cam]
    prm.min_x, prm.max_x, prm.min_y, prm.max_y = [float(v) for v in bounds]
    sf = np.asarray(scale_factors, np.float32)
    for l in range(len(sf)):
        prm.scale_factors[l] = float(sf[l])
    prm.nlevels = len(sf); prm.mbf = float(mbf); prm.mb = float(mb); prm.th = float(th)
    prm.mono = 1 if mono else 0; prm.check_orientation = 1 if check_orientation else 0
    match = np.full((P, cap_c), -1, np.int32); nm = np.zeros(P, np.int32)
    _check(lib().orbx_search_by_projection(kl.ctypes.data_as(_vp), _ptr(fl), wd.ctypes.data_as(_vp), _ptr(md), _ptr(nl, _i32p),
                                           kc.ctypes.data_as(_vp), _ptr(dc), ur.ctypes.data_as(_vp) if has_ur else None,
                                           tk.ctypes.data_as(_vp) if has_tk else None, _ptr(nc, _i32p), poses.ctypes.data_as(_vp), P, cap_l,
                                           cap_c, C.byref(prm), _ptr(match, _i32p), _ptr(nm, _i32p), device))
    return [(int(nm[p]), match[p, :nc[p]].copy()) for p in range(P)]


def search_by_projection_map(cases, bounds, scale_factors, th=3.0, far_points=False, th_far_points=50.0, nnratio=0.8, device=-1):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints) (ORBmatcher.cc:43-210, Tracking::SearchLocalPoints) on the
    device for a list of independent (map points, frame) problems.  Each case: track (n x 5), level, flags, mp_desc, kps_cur, desc_cur and
    optionally cur_u_right, cur_taken.  -> [(nmatches, match_cur[n_cur])] per case."""
    P = len(cases)
    cap_m = max([len(c["level"]) for c in cases] + [1])
    cap_c = max([len(c["kps_cur"]) for c in cases] + [1])
    tr = np.zeros((P, cap_m, 5), np.float32); lv = np.zeros((P, cap_m), np.int32); fl = np.zeros((P, cap_m), np.uint8)
    md = np.zeros((P, cap_m, 32), np.uint8); kc = np.zeros((P, cap_c), dtype=KP_DTYPE); dc = np.zeros((P, cap_c, 32), np.uint8)
    has_ur = cases[0].get("cur_u_right") is not None
    has_tk = cases[0].get("cur_taken") is not None
    ur = np.full((P, cap_c), -1, np.float32); tk = np.zeros((P, cap_c), np.uint8)
    nm_ = np.zeros(P, np.int32); nc = np.zeros(P, np.int32)
    for p, c in enumerate(cases):
        a, b = len(c["level"]), len(c["kps_cur"])
        nm_[p], nc[p] = a, b
        tr[p, :a] = np.asarray(c["track"], np.float32).reshape(-1, 5); lv[p, :a] = c["level"]; fl[p, :a] = c["flags"]; md[p, :a] = c["mp_desc"]
        kc[p, :b] = c["kps_cur"]; dc[p, :b] = c["desc_cur"]
        if has_ur: ur[p, :b] = c["cur_u_right"]
        if has_tk: tk[p, :b] = c["cur_taken"]
    prm = ProjParams()
    prm.min_x, prm.max_x, prm.min_y, prm.max_y = [float(v) for v in bounds]
    sf = np.asarray(scale_factors, np.float32)
    for l in range(len(sf)):
        prm.scale_factors[l] = float(sf[l])
    prm.nlevels = len(sf); prm.th = float(th)
    match = np.full((P, cap_c), -1, np.int32); nmat = np.zeros(P, np.int32)
    _check(lib().orbx_search_by_projection_map(tr.ctypes.data_as(_vp), _ptr(lv, _i32p), _ptr(fl), _ptr(md), _ptr(nm_, _i32p),
                                               kc.ctypes.data_as(_vp), _ptr(dc), ur.ctypes.data_as(_vp) if has_ur else None,
                                               tk.ctypes.data_as(_vp) if has_tk else None, _ptr(nc, _i32p), P, cap_m, cap_c, C.byref(prm),
                                               float(nnratio), 1 if far_points else 0, float(th_far_points), _ptr(match, _i32p),
                                               _ptr(nmat, _i32p), device))
    return [(int(nmat[p]), match[p, :nc[p]].copy()) for p in range(P)]


def accept_bow(d1, d2, nnratio=0.7, th=TH_LOW):
    """The 'BoW KF->F' accept rule (ORBmatcher.cc:327-329): best1 <= TH_LOW and (float)best1 < nnratio*(float)best2."""
    d1 = np.asarray(d1)
    d2 = np.asarray(d2)
    return (d1 <= th) & (d1.astype(np.float32) < np.float32(nnratio) * d2.astype(np.float32))


class ORBVocabulary:
    """Device-resident DBoW2 vocabulary tree (ORB_SLAM3/include/ORBVocabulary.h:27-28) given as flat arrays; transform() is
    Frame::ComputeBoW (src/Frame.cc:738-745) = TemplatedVocabulary::transform(features, BowVector, FeatureVector, levelsup)."""

    def __init__(self, parent, node_desc, weight, L, device=-1):
        parent = np.ascontiguousarray(parent, np.int32)
        node_desc = np.ascontiguousarray(node_desc, np.uint8).reshape(-1, 32)
        weight = np.ascontiguousarray(weight, np.float64)
        if not (len(parent) == len(node_desc) == len(weight)):
            raise ValueError("parent, node_desc and weight must have one entry per node")
        self._h = _vp()
        _check(lib().orbx_voc_create(len(parent), _ptr(parent, _i32p), _ptr(node_desc), weight.ctypes.data_as(C.POINTER(C.c_double)),
                                     int(L), device, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None):
            lib().orbx_voc_destroy(self._h)
            self._h = None

    __del__ = close

    def info(self):
        v = [np.zeros(1, np.int32) for _ in range(4)]
        _check(lib().orbx_voc_info(self._h, *[_ptr(a, _i32p) for a in v]))
        return dict(n_nodes=int(v[0][0]), n_words=int(v[1][0]), max_children=int(v[2][0]), L=int(v[3][0]))

    @staticmethod
    def _outputs(F, cap):
        return (np.zeros((F, cap), np.int32), np.zeros((F, cap), np.float64), np.zeros(F, np.int32), np.zeros((F, cap), np.int32),
                np.zeros((F, cap + 1), np.int32), np.zeros((F, cap), np.int32), np.zeros(F, np.int32))

    @staticmethod
    def _unpack(o):
        wi, wv, nw, ni, nb, fi, nn = o
        out = []
        for f in range(len(nw)):
            a, b = int(nw[f]), int(nn[f])
            out.append(dict(word_ids=wi[f, :a].copy(), word_vals=wv[f, :a].copy(), node_ids=ni[f, :b].copy(),
                            node_begin=nb[f, :b + 1].copy(), feat_idx=fi[f, :int(nb[f, b])].copy()))
        return out

    def transform(self, desc, levelsup=4):
        """desc: (n, 32) descriptors of ONE frame -> dict(word_ids, word_vals, node_ids, node_begin, feat_idx)."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        return self.transform_batch([desc], levelsup)[0]

    def transform_batch(self, descs, levelsup=4):
        F = len(descs)
        n = np.asarray([len(d) for d in descs], np.int32)
        cap = max(int(n.max()) if F else 0, 1)
        buf = np.zeros((F, cap, 32), np.uint8)
        for f, d in enumerate(descs):
            buf[f, :len(d)] = np.asarray(d, np.uint8).reshape(-1, 32)
        o = self._outputs(F, cap)
        _check(lib().orbx_bow_transform(self._h, _ptr(buf), _ptr(n, _i32p), F, cap, levelsup, *[a.ctypes.data for a in o]))
        return self._unpack(o)

    def transform_extracted(self, extractor, levelsup=4, raw=None):
        """BoW / feature vectors of every frame of `extractor`'s last extraction, descriptors read in place on the device.
        raw = the tuple of preallocated output arrays of a previous call's `outputs_for(extractor)`: filled and returned
        as they are (the C ABI's own layout, no per-frame unpacking)."""
        o = raw if raw is not None else self.outputs_for(extractor)
        _check(lib().orbx_bow_transform_extracted(self._h, extractor._h, levelsup, *[a.ctypes.data for a in o]))
        return o if raw is not None else self._unpack(o)

    def outputs_for(self, extractor):
        return self._outputs(extractor._frames, extractor.results_device()[4])


def search_by_bow(desc_kf, valid_kf, angle_kf, fv_kf, desc_f, angle_f, fv_f, nnratio=0.7, check_orientation=True, device=-1):
    """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) (ORBmatcher.cc:223-425, monocular) -> (nmatches, match_f): match_f[j] =
    keyframe feature whose MapPoint goes to frame feature j, or -1.  fv_kf / fv_f: FeatureVector dicts of ORBVocabulary.transform."""
    desc_kf = np.ascontiguousarray(desc_kf, np.uint8).reshape(-1, 32)
    desc_f = np.ascontiguousarray(desc_f, np.uint8).reshape(-1, 32)
    valid_kf = np.ascontiguousarray(valid_kf, np.uint8)
    angle_kf = np.ascontiguousarray(angle_kf, np.float32)
    angle_f = np.ascontiguousarray(angle_f, np.float32)
    a = [np.ascontiguousarray(fv_kf[k], np.int32) for k in ("node_ids", "node_begin", "feat_idx")]
    b = [np.ascontiguousarray(fv_f[k], np.int32) for k in ("node_ids", "node_begin", "feat_idx")]
    out = np.full(max(len(desc_f), 1), -1, np.int32)
    n = np.zeros(1, np.int32)
    _check(lib().orbx_search_by_bow(_ptr(desc_kf), len(desc_kf), _ptr(valid_kf), _ptr(angle_kf, _f32p), *[_ptr(x, _i32p) for x in a],
                                    len(a[0]), _ptr(desc_f), len(desc_f), _ptr(angle_f, _f32p), *[_ptr(x, _i32p) for x in b], len(b[0]),
                                    float(nnratio), int(bool(check_orientation)), _ptr(out, _i32p), _ptr(n, _i32p), device))
    return int(n[0]), out[:len(desc_f)]


def search_by_bow_kf(desc1, valid1, angle1, fv1, desc2, valid2, angle2, fv2, nnratio=0.75, check_orientation=True, device=-1):
    """ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, ...) (ORBmatcher.cc:765-905, monocular) -> (nmatches, match12): match12[i] =
    feature of keyframe 2 whose MapPoint goes to feature i of keyframe 1, or -1."""
    desc1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
    desc2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
    valid1 = np.ascontiguousarray(valid1, np.uint8)
    valid2 = np.ascontiguousarray(valid2, np.uint8)
    angle1 = np.ascontiguousarray(angle1, np.float32)
    angle2 = np.ascontiguousarray(angle2, np.float32)
    a = [np.ascontiguousarray(fv1[k], np.int32) for k in ("node_ids", "node_begin", "feat_idx")]
    b = [np.ascontiguousarray(fv2[k], np.int32) for k in ("node_ids", "node_begin", "feat_idx")]
    out = np.full(max(len(desc1), 1), -1, np.int32)
    n = np.zeros(1, np.int32)
    _check(lib().orbx_search_by_bow_kf(_ptr(desc1), len(desc1), _ptr(valid1), _ptr(angle1, _f32p), *[_ptr(x, _i32p) for x in a], len(a[0]),
                                       _ptr(desc2), len(desc2), _ptr(valid2), _ptr(angle2, _f32p), *[_ptr(x, _i32p) for x in b], len(b[0]),
                                       float(nnratio), int(bool(check_orientation)), _ptr(out, _i32p), _ptr(n, _i32p), device))
    return int(n[0]), out[:len(desc1)]


class ShardedMatcher:
    """One rank of the sharded top-2 matcher (config 5): local top-2 -> ncclAllGather -> merge on every rank."""

    def __init__(self, rank, nranks, unique_id, device):
        self._c = _vp()
        uid = np.frombuffer(bytes(unique_id), dtype=np.uint8).copy()
        assert uid.size == 128
        _check(lib().orbx_comm_init(rank, nranks, _ptr(uid), device, C.byref(self._c)))
        self.rank, self.nranks = rank, nranks

    @staticmethod
    def unique_id():
        uid = np.zeros(128, dtype=np.uint8)
        _check(lib().orbx_comm_unique_id(_ptr(uid)))
        return uid.tobytes()

    def match(self, d_q, nq, d_db_local, nd_local, index_base, d_idx1, d_d1, d_idx2, d_d2):
        _check(lib().orbx_match_top2_sharded(self._c, d_q, nq, d_db_local, nd_local, index_base, d_idx1, d_d1, d_idx2,
                                             d_d2))

    def match_async(self, d_q, nq, d_db_local, nd_local, index_base, d_idx1, d_d1, d_idx2, d_d2):
        """Enqueue only (orbx_match_top2_sharded_async); sync() waits for every call queued so far."""
        _check(lib().orbx_match_top2_sharded_async(self._c, d_q, nq, d_db_local, nd_local, index_base, d_idx1, d_d1, d_idx2, d_d2))

    def sync(self):
        _check(lib().orbx_comm_sync(self._c))

    def close(self):
        if getattr(self, "_c", None) is not None and self._c.value:
            lib().orbx_comm_destroy(self._c)
            self._c = _vp()

    __del__ = close


def shard_range(n, rank, nranks):
    """Rows [lo, hi) of an n-row database owned by `rank` (rows [r*N/G, (r+1)*N/G))."""
    return (n * rank) // nranks, (n * (rank + 1)) // nranks


def merge_top2_host(keys1, keys2):
    """Host restatement of the merge rule for CPU-side (gloo) tests of the exchange step: keys are
    distance << 40 | global row, sets along axis 0; returns merged (k1, k2)."""
    allk = np.concatenate([np.asarray(keys1, dtype=np.uint64), np.asarray(keys2, dtype=np.uint64)], axis=0)
    s = np.sort(allk, axis=0)
    return s[0], s[1]
