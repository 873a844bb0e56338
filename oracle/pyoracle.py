"""ctypes bindings for the oracle libraries — TEST INFRASTRUCTURE, not product code.

  RefExtractor   the reference's own ORBextractor.cc compiled unchanged on the minicv shim
                 (oracle/_ref/liborbref_{strict,native}.so, built by oracle/Makefile)
  port           oracle/liborb_oracle.so: CPU restatement of the path + minicv primitives
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

_u8p = C.POINTER(C.c_uint8)


def _ptr(a, t=_u8p):
    return a.ctypes.data_as(t)


def build(quiet=True):
    """(Re)build the oracle libraries; the _ref targets only where /root/reference exists."""
    r = subprocess.run(["make", "-C", HERE, "all"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    if not quiet:
        print(r.stdout)


def ref_path(mode="strict"):
    return os.path.join(HERE, "_ref", f"liborbref_{mode}.so")


def have_ref(mode="strict"):
    return os.path.exists(ref_path(mode))


_ref_libs = {}


def _ref_lib(mode):
    if mode not in _ref_libs:
        lib = C.CDLL(ref_path(mode))
        lib.orbref_create.restype = C.c_void_p
        lib.orbref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        lib.orbref_destroy.argtypes = [C.c_void_p]
        lib.orbref_extract.restype = C.c_int
        lib.orbref_extract.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                       _u8p, C.c_int, C.POINTER(C.c_int)]
        lib.orbref_pyramid.restype = C.c_int
        lib.orbref_pyramid.argtypes = [C.c_void_p, C.c_int, _u8p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        lib.orbref_tables.restype = C.c_int
        lib.orbref_tables.argtypes = [C.c_void_p] + [C.POINTER(C.c_float)] * 4 + [C.POINTER(C.c_int)] * 2
        lib.orbref_scale_factor.restype = C.c_float
        lib.orbref_scale_factor.argtypes = [C.c_void_p]
        _ref_libs[mode] = lib
    return _ref_libs[mode]


class RefExtractor:
    """The reference ORB_SLAM3::ORBextractor (ORBextractor.h:43-109) behind ref_wrap.cpp."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, mode="strict"):
        self.lib = _ref_lib(mode)
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self.h = self.lib.orbref_create(nfeatures, scale_factor, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.orbref_destroy(self.h)
            self.h = None

    def extract(self, img, lap=(0, 1000)):
        """-> (mono_index, keypoints[KP_DTYPE], descriptors[n,32])"""
        cap = self.nfeatures + 4 * self.nlevels + 64
        kps = np.zeros(cap, dtype=KP_DTYPE)
        desc = np.zeros((cap, 32), dtype=np.uint8)
        n = C.c_int(0)
        if img is None or img.size == 0:
            mono = self.lib.orbref_extract(self.h, None, 0, 0, 0, lap[0], lap[1], None, None, 0, C.byref(n))
            return mono, kps[:0], desc[:0]
        img = np.ascontiguousarray(img, dtype=np.uint8)
        h, w = img.shape
        mono = self.lib.orbref_extract(self.h, _ptr(img), w, h, w, lap[0], lap[1], kps.ctypes.data_as(C.c_void_p),
                                       _ptr(desc), cap, C.byref(n))
        assert n.value <= cap
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def pyramid(self, level):
        """Padded plane (h+38, w+38) of the last extract() call."""
        w, h = C.c_int(0), C.c_int(0)
        if self.lib.orbref_pyramid(self.h, level, None, C.byref(w), C.byref(h)) != 0:
            return None
        out = np.zeros((h.value + 38, w.value + 38), dtype=np.uint8)
        self.lib.orbref_pyramid(self.h, level, _ptr(out), None, None)
        return out

    def tables(self):
        nl = self.nlevels
        f = [np.zeros(nl, dtype=np.float32) for _ in range(4)]
        nf = np.zeros(nl, dtype=np.int32)
        um = np.zeros(16, dtype=np.int32)
        self.lib.orbref_tables(self.h, *[_ptr(a, C.POINTER(C.c_float)) for a in f], _ptr(nf, C.POINTER(C.c_int)),
                               _ptr(um, C.POINTER(C.c_int)))
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3], nfeatures=nf, umax=um,
                    scale_factor=self.lib.orbref_scale_factor(self.h))


# --------------------------------------------------------------------------------------------
# the port oracle (oracle/liborb_oracle.so)
# --------------------------------------------------------------------------------------------
_i32p = C.POINTER(C.c_int32)
_f32p = C.POINTER(C.c_float)
_port = None


def port_path():
    return os.path.join(HERE, "liborb_oracle.so")


def port_lib():
    global _port
    if _port is None:
        if not os.path.exists(port_path()):
            build()
        lib = C.CDLL(port_path())
        lib.orbo_create.restype = C.c_void_p
        lib.orbo_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int]
        lib.orbo_destroy.argtypes = [C.c_void_p]
        lib.orbo_extract.restype = C.c_int
        lib.orbo_extract.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, _u8p,
                                     C.c_int, C.POINTER(C.c_int)]
        lib.orbo_pyramid.restype = C.c_int
        lib.orbo_pyramid.argtypes = [C.c_void_p, C.c_int, _u8p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        lib.orbo_tables.restype = C.c_int
        lib.orbo_tables.argtypes = [C.c_void_p] + [_f32p] * 4 + [C.POINTER(C.c_int)] * 2
        lib.orbo_candidates.restype = C.c_int
        lib.orbo_candidates.argtypes = [C.c_void_p, C.c_int, _i32p, C.c_int]
        lib.orbo_level_keypoints.restype = C.c_int
        lib.orbo_level_keypoints.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        lib.orbo_quadtree.restype = C.c_int
        lib.orbo_quadtree.argtypes = [_i32p, C.c_int, C.c_int, C.c_int, C.c_int, _i32p, C.c_int]
        lib.orbo_sort_check.argtypes = [C.POINTER(C.c_int64), C.c_int, _i32p, _i32p]
        lib.orbo_rbrief.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, _u8p]
        lib.orbo_sincosf.argtypes = [_f32p, _f32p, _f32p, C.c_int]
        lib.orbo_hamming.restype = C.c_int
        lib.orbo_hamming.argtypes = [_u8p, _u8p]
        lib.orbo_match_top2.argtypes = [_u8p, C.c_int, _u8p, C.c_long, _i32p, _i32p, _i32p, _i32p, C.c_int]
        lib.orbo_match_stereo_banded.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_void_p, _u8p, C.c_int, _f32p, C.c_int,
                                                 C.c_float, _i32p, _i32p]
        lib.orbo_accept_top2.argtypes = [C.c_int, C.c_float, _i32p, _i32p, C.POINTER(C.c_int64), _i32p, _i32p, C.c_int, _u8p]
        lib.orbo_search_grid_top2.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, _u8p, _f32p,
                                              C.c_void_p, _u8p, C.c_int] + [_i32p] * 6
        lib.orbo_compute_stereo_matches.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_void_p, _u8p, C.c_int, _f32p, _f32p, C.c_int,
                                                    C.POINTER(C.c_void_p), _i32p, C.POINTER(C.c_void_p), _i32p, _i32p, C.c_int,
                                                    C.c_float, C.c_float, _f32p, _f32p]
        lib.orbo_extract_many.restype = C.c_long
        lib.orbo_extract_many.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, _u8p, C.c_int,
                                          C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        lib.minicv_resize_linear.argtypes = [_u8p, C.c_int, C.c_int, _u8p, C.c_int, C.c_int]
        lib.minicv_border101.argtypes = [_u8p, C.c_int, C.c_int, _u8p, C.c_int]
        lib.minicv_fast.restype = C.c_int
        lib.minicv_fast.argtypes = [_u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _i32p, C.c_int]
        lib.minicv_gauss7.argtypes = [_u8p, C.c_int, C.c_int, _u8p]
        lib.minicv_fast_atan2.argtypes = [_f32p, _f32p, _f32p, C.c_int]
        lib.minicv_undistort_points.argtypes = [_f32p, C.c_int, _f32p, _f32p, C.c_int, _f32p, _f32p]
        lib.orbo_undistort_keypoints.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, C.c_int, _f32p, C.c_void_p]
        lib.orbo_image_bounds.argtypes = [C.c_int, C.c_int, _f32p, _f32p, C.c_int, _f32p, _f32p]
        _port = lib
    return _port


class PortExtractor:
    """The restated extractor (orb_oracle.cpp), same call shape as RefExtractor."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, mode="strict"):
        self.lib = port_lib()
        self.nlevels, self.nfeatures = nlevels, nfeatures
        self.h = self.lib.orbo_create(nfeatures, scale_factor, nlevels, ini_th, min_th, 0 if mode == "strict" else 1)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.orbo_destroy(self.h)
            self.h = None

    def extract(self, img, lap=(0, 1000)):
        cap = self.nfeatures + 8 * self.nlevels + 64
        kps = np.zeros(cap, dtype=KP_DTYPE)
        desc = np.zeros((cap, 32), dtype=np.uint8)
        n = C.c_int(0)
        if img is None or img.size == 0:
            return self.lib.orbo_extract(self.h, None, 0, 0, 0, lap[0], lap[1], None, None, 0, C.byref(n)), kps[:0], desc[:0]
        img = np.ascontiguousarray(img, dtype=np.uint8)
        h, w = img.shape
        mono = self.lib.orbo_extract(self.h, _ptr(img), w, h, w, lap[0], lap[1], kps.ctypes.data_as(C.c_void_p),
                                     _ptr(desc), cap, C.byref(n))
        assert mono >= -1, "oracle keypoint buffer too small"
        return mono, kps[:n.value].copy(), desc[:n.value].copy()

    def pyramid(self, level):
        w, h = C.c_int(0), C.c_int(0)
        if self.lib.orbo_pyramid(self.h, level, None, C.byref(w), C.byref(h)) != 0:
            return None
        out = np.zeros((h.value + 38, w.value + 38), dtype=np.uint8)
        self.lib.orbo_pyramid(self.h, level, _ptr(out), None, None)
        return out

    def candidates(self, level):
        n = self.lib.orbo_candidates(self.h, level, None, 0)
        out = np.zeros((max(n, 1), 3), dtype=np.int32)
        self.lib.orbo_candidates(self.h, level, _ptr(out, _i32p), n)
        return out[:n]

    def level_keypoints(self, level):
        n = self.lib.orbo_level_keypoints(self.h, level, None, 0)
        out = np.zeros(max(n, 1), dtype=KP_DTYPE)
        self.lib.orbo_level_keypoints(self.h, level, out.ctypes.data_as(C.c_void_p), n)
        return out[:n]

    def tables(self):
        nl = self.nlevels
        f = [np.zeros(nl, dtype=np.float32) for _ in range(4)]
        nf = np.zeros(nl, dtype=np.int32)
        um = np.zeros(16, dtype=np.int32)
        self.lib.orbo_tables(self.h, *[_ptr(a, _f32p) for a in f], _ptr(nf, C.POINTER(C.c_int)), _ptr(um, C.POINTER(C.c_int)))
        return dict(scale=f[0], inv_scale=f[1], sigma2=f[2], inv_sigma2=f[3], nfeatures=nf, umax=um)


def quadtree(cand, width, height, N):
    cand = np.ascontiguousarray(cand, dtype=np.int32).reshape(-1, 3)
    keep = np.zeros(N + 8, dtype=np.int32)
    n = port_lib().orbo_quadtree(_ptr(cand, _i32p), len(cand), width, height, N, _ptr(keep, _i32p), len(keep))
    assert n <= len(keep)
    return keep[:n].copy()


def sort_check(keys):
    keys = np.ascontiguousarray(keys, dtype=np.int64)
    a = np.zeros(len(keys), dtype=np.int32)
    b = np.zeros(len(keys), dtype=np.int32)
    port_lib().orbo_sort_check(keys.ctypes.data_as(C.POINTER(C.c_int64)), len(keys), _ptr(a, _i32p), _ptr(b, _i32p))
    return a, b


def hamming(a, b):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    b = np.ascontiguousarray(b, dtype=np.uint8)
    return port_lib().orbo_hamming(_ptr(a), _ptr(b))


def match_top2(q, db, threads=1):
    q = np.ascontiguousarray(q, dtype=np.uint8).reshape(-1, 32)
    db = np.ascontiguousarray(db, dtype=np.uint8).reshape(-1, 32)
    out = [np.zeros(len(q), dtype=np.int32) for _ in range(4)]
    port_lib().orbo_match_top2(_ptr(q), len(q), _ptr(db), len(db), *[_ptr(o, _i32p) for o in out], threads)
    return out  # idx1, d1, idx2, d2


def match_stereo_banded(kl, dl, kr, dr, scale_factors, n_rows, max_d):
    kl = np.ascontiguousarray(kl); kr = np.ascontiguousarray(kr)
    dl = np.ascontiguousarray(dl, dtype=np.uint8); dr = np.ascontiguousarray(dr, dtype=np.uint8)
    sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
    bi = np.zeros(len(kl), dtype=np.int32); bd = np.zeros(len(kl), dtype=np.int32)
    port_lib().orbo_match_stereo_banded(kl.ctypes.data_as(C.c_void_p), _ptr(dl), len(kl), kr.ctypes.data_as(C.c_void_p),
                                        _ptr(dr), len(kr), _ptr(sf, _f32p), n_rows, max_d, _ptr(bi, _i32p), _ptr(bd, _i32p))
    return bi, bd


def accept_top2(rule, d1, d2, nnratio=0.7, idx2=None, level1=None, level2=None):
    """ORBmatcher's accept rules (see include/orbx.h orbx_accept_rule) -> bool array."""
    d1 = np.ascontiguousarray(d1, dtype=np.int32); d2 = np.ascontiguousarray(d2, dtype=np.int32)
    keep = []

    def opt(a, dt, t):
        if a is None:
            return None
        a = np.ascontiguousarray(a, dtype=dt)
        keep.append(a)
        return _ptr(a, t)
    acc = np.zeros(max(len(d1), 1), dtype=np.uint8)
    port_lib().orbo_accept_top2(int(rule), float(nnratio), _ptr(d1, _i32p), _ptr(d2, _i32p), opt(idx2, np.int64, C.POINTER(C.c_int64)),
                                opt(level1, np.int32, _i32p), opt(level2, np.int32, _i32p), len(d1), _ptr(acc))
    return acc[:len(d1)].astype(bool)


def search_grid_top2(kps_t, desc_t, bounds, queries, desc_q, skip_t=None, u_right_t=None):
    """Frame::GetFeaturesInArea + best/second loop -> idx1, d1, idx2, d2, lvl1, lvl2."""
    kps_t = np.ascontiguousarray(kps_t); queries = np.ascontiguousarray(queries)
    desc_t = np.ascontiguousarray(desc_t, dtype=np.uint8); desc_q = np.ascontiguousarray(desc_q, dtype=np.uint8)
    assert kps_t.dtype.itemsize == 28 and queries.dtype.itemsize == 24
    sk = None if skip_t is None else np.ascontiguousarray(skip_t, dtype=np.uint8)
    ur = None if u_right_t is None else np.ascontiguousarray(u_right_t, dtype=np.float32)
    nq = len(queries)
    out = [np.zeros(max(nq, 1), dtype=np.int32) for _ in range(6)]
    port_lib().orbo_search_grid_top2(kps_t.ctypes.data_as(C.c_void_p), _ptr(desc_t), len(kps_t), *[float(b) for b in bounds],
                                     None if sk is None else _ptr(sk), None if ur is None else _ptr(ur, _f32p),
                                     queries.ctypes.data_as(C.c_void_p), _ptr(desc_q), nq, *[_ptr(o, _i32p) for o in out])
    return [o[:nq] for o in out]


def compute_stereo_matches(kl, dl, kr, dr, scale_factors, inv_scale_factors, pyr_l, pyr_r, bf, b):
    """Frame::ComputeStereoMatches (Frame.cc:811-980).  pyr_l / pyr_r: lists of PADDED planes (h+38, w+38) per level,
    as PortExtractor.pyramid returns them -> (mvuRight, mvDepth)."""
    kl = np.ascontiguousarray(kl); kr = np.ascontiguousarray(kr)
    dl = np.ascontiguousarray(dl, dtype=np.uint8); dr = np.ascontiguousarray(dr, dtype=np.uint8)
    sf = np.ascontiguousarray(scale_factors, dtype=np.float32); isf = np.ascontiguousarray(inv_scale_factors, dtype=np.float32)
    nl = len(sf)
    pl = [np.ascontiguousarray(p, dtype=np.uint8) for p in pyr_l]
    pr = [np.ascontiguousarray(p, dtype=np.uint8) for p in pyr_r]
    ptr_l = (C.c_void_p * nl)(*[p.ctypes.data + 19 * p.shape[1] + 19 for p in pl])
    ptr_r = (C.c_void_p * nl)(*[p.ctypes.data + 19 * p.shape[1] + 19 for p in pr])
    st_l = np.array([p.shape[1] for p in pl], dtype=np.int32); st_r = np.array([p.shape[1] for p in pr], dtype=np.int32)
    cols = np.array([p.shape[1] - 38 for p in pl], dtype=np.int32)
    n_rows = pl[0].shape[0] - 38
    n = len(kl)
    ur = np.zeros(max(n, 1), dtype=np.float32); dp = np.zeros(max(n, 1), dtype=np.float32)
    port_lib().orbo_compute_stereo_matches(kl.ctypes.data_as(C.c_void_p), _ptr(dl), n, kr.ctypes.data_as(C.c_void_p), _ptr(dr),
                                           len(kr), _ptr(sf, _f32p), _ptr(isf, _f32p), nl, ptr_l, _ptr(st_l, _i32p), ptr_r,
                                           _ptr(st_r, _i32p), _ptr(cols, _i32p), n_rows, float(bf), float(b), _ptr(ur, _f32p),
                                           _ptr(dp, _f32p))
    return ur[:n], dp[:n]


def _cam(K, dist, P):
    K = np.ascontiguousarray(K, dtype=np.float32).reshape(3, 3)
    P = K if P is None else np.ascontiguousarray(P, dtype=np.float32).reshape(3, 3)
    return K, np.ascontiguousarray(dist, dtype=np.float32).ravel(), P


def undistort_points(pts, K, dist, P=None):
    """minicv's cv::undistortPoints(pts, K, dist, R = I, P) -> n x 2 float32."""
    K, dist, P = _cam(K, dist, P)
    pts = np.ascontiguousarray(pts, dtype=np.float32).reshape(-1, 2)
    out = np.zeros_like(pts)
    port_lib().minicv_undistort_points(_ptr(pts, _f32p), len(pts), _ptr(K, _f32p), _ptr(dist, _f32p), len(dist), _ptr(P, _f32p),
                                       _ptr(out, _f32p))
    return out


def _undistort_keypoints(fn, kps, K, dist, P):
    K, dist, P = _cam(K, dist, P)
    kps = np.ascontiguousarray(kps)
    assert kps.dtype.itemsize == 28
    out = np.zeros(max(len(kps), 1), dtype=kps.dtype)
    fn(kps.ctypes.data_as(C.c_void_p), len(kps), _ptr(K, _f32p), _ptr(dist, _f32p), len(dist), _ptr(P, _f32p),
       out.ctypes.data_as(C.c_void_p))
    return out[:len(kps)]


def _image_bounds(fn, w, h, K, dist, P):
    K, dist, P = _cam(K, dist, P)
    out = np.zeros(4, np.float32)
    fn(int(w), int(h), _ptr(K, _f32p), _ptr(dist, _f32p), len(dist), _ptr(P, _f32p), _ptr(out, _f32p))
    return out


def undistort_keypoints(kps, K, dist, P=None):
    """Frame::UndistortKeyPoints (Frame.cc:747-780): mvKeys -> mvKeysUn."""
    return _undistort_keypoints(port_lib().orbo_undistort_keypoints, kps, K, dist, P)


def image_bounds(w, h, K, dist, P=None):
    """Frame::ComputeImageBounds (Frame.cc:782-809) -> (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
    return _image_bounds(port_lib().orbo_image_bounds, w, h, K, dist, P)


def ref_undistort_keypoints(kps, K, dist, P=None):
    """Frame::UndistortKeyPoints, the reference's own body (librefslices.so)."""
    return _undistort_keypoints(_slices().refslice_undistort_keypoints, kps, K, dist, P)


def ref_image_bounds(w, h, K, dist, P=None):
    """Frame::ComputeImageBounds, the reference's own body (librefslices.so)."""
    return _image_bounds(_slices().refslice_image_bounds, w, h, K, dist, P)


def _search_for_initialization(fn, kps1, desc1, kps2, desc2, bounds, prev_matched, window_size, nnratio, check_orientation):
    kps1 = np.ascontiguousarray(kps1); kps2 = np.ascontiguousarray(kps2)
    assert kps1.dtype.itemsize == 28 and kps2.dtype.itemsize == 28
    desc1 = np.ascontiguousarray(desc1, dtype=np.uint8); desc2 = np.ascontiguousarray(desc2, dtype=np.uint8)
    prev = np.array(prev_matched, dtype=np.float32).reshape(-1, 2).copy()
    assert len(prev) == len(kps1)
    m12 = np.zeros(max(len(kps1), 1), dtype=np.int32)
    fn.restype = C.c_int
    n = fn(kps1.ctypes.data_as(C.c_void_p), _ptr(desc1), len(kps1), kps2.ctypes.data_as(C.c_void_p), _ptr(desc2), len(kps2),
           *[C.c_float(float(b)) for b in bounds], prev.ctypes.data_as(_f32p), C.c_int(int(window_size)), C.c_float(float(nnratio)),
           C.c_int(int(bool(check_orientation))), _ptr(m12, _i32p))
    return n, m12[:len(kps1)], prev


def search_for_initialization(kps1, desc1, kps2, desc2, bounds, prev_matched, window_size=100, nnratio=0.9, check_orientation=True):
    """ORBmatcher::SearchForInitialization (ORBmatcher.cc:648-763) -> (nmatches, vnMatches12, vbPrevMatched after the call)."""
    return _search_for_initialization(port_lib().orbo_search_for_initialization, kps1, desc1, kps2, desc2, bounds, prev_matched,
                                      window_size, nnratio, check_orientation)


def ref_search_for_initialization(kps1, desc1, kps2, desc2, bounds, prev_matched, window_size=100, nnratio=0.9,
                                  check_orientation=True):
    """The same through the reference's own body (librefslices.so)."""
    return _search_for_initialization(_slices().refslice_search_for_initialization, kps1, desc1, kps2, desc2, bounds, prev_matched,
                                      window_size, nnratio, check_orientation)


def _search_by_projection_ff(fn, kps_last, last_flags, world, mp_desc, kps_cur, desc_cur, pose_cw, pose_lw, cam, bounds, scale_factors,
                             th, mono, mbf=0.0, mb=0.0, cur_u_right=None, cur_taken=None, nnratio=0.9, check_orientation=True):
    kps_last = np.ascontiguousarray(kps_last); kps_cur = np.ascontiguousarray(kps_cur)
    assert kps_last.dtype.itemsize == 28 and kps_cur.dtype.itemsize == 28
    n_last, n_cur = len(kps_last), len(kps_cur)
    last_flags = np.ascontiguousarray(last_flags, dtype=np.uint8)
    world = np.ascontiguousarray(world, dtype=np.float32).reshape(-1, 3)
    mp_desc = np.ascontiguousarray(mp_desc, dtype=np.uint8); desc_cur = np.ascontiguousarray(desc_cur, dtype=np.uint8)
    pose_cw = np.ascontiguousarray(pose_cw, dtype=np.float32); pose_lw = np.ascontiguousarray(pose_lw, dtype=np.float32)
    cam = np.ascontiguousarray(cam, dtype=np.float32); sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
    ur = None if cur_u_right is None else np.ascontiguousarray(cur_u_right, dtype=np.float32)
    tk = None if cur_taken is None else np.ascontiguousarray(cur_taken, dtype=np.uint8)
    out = np.full(max(n_cur, 1), -1, dtype=np.int32)
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, _u8p, _f32p, _u8p, C.c_int, C.c_void_p, _u8p, _f32p, _u8p, C.c_int, _f32p, _f32p, _f32p, C.c_float,
                   C.c_float, C.c_float, C.c_float, _f32p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_int, C.c_float, C.c_int, _i32p]
    n = fn(kps_last.ctypes.data_as(C.c_void_p), _ptr(last_flags), _ptr(world, _f32p), _ptr(mp_desc), n_last,
           kps_cur.ctypes.data_as(C.c_void_p), _ptr(desc_cur), None if ur is None else _ptr(ur, _f32p), None if tk is None else _ptr(tk),
           n_cur, _ptr(pose_cw, _f32p), _ptr(pose_lw, _f32p), _ptr(cam, _f32p), *[float(b) for b in bounds], _ptr(sf, _f32p), len(sf),
           float(mbf), float(mb), float(th), 1 if mono else 0, float(nnratio), 1 if check_orientation else 0, _ptr(out, _i32p))
    return int(n), out[:n_cur]


def search_by_projection_ff(*a, **k):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (ORBmatcher.cc:1676-1887), Nleft == -1 frames ->
    (nmatches, match_cur[n_cur]: index of the last-frame feature whose map point the current feature holds, or -1).
    poses = (qx, qy, qz, qw, tx, ty, tz); cam = (fx, fy, cx, cy); last_flags bit 0 = map point and not outlier, bit 1 = Observations() > 0."""
    return _search_by_projection_ff(port_lib().orbo_search_by_projection_ff, *a, **k)


def ref_search_by_projection_ff(*a, **k):
    """The same through the reference's own body (librefslices.so)."""
    return _search_by_projection_ff(_slices().refslice_search_by_projection_ff, *a, **k)


def _search_by_projection_map(fn, track, level, flags, mp_desc, kps_cur, desc_cur, bounds, scale_factors, th=3.0, far_points=False,
                              th_far_points=50.0, nnratio=0.8, cur_u_right=None, cur_taken=None):
    track = np.ascontiguousarray(track, dtype=np.float32).reshape(-1, 5)
    level = np.ascontiguousarray(level, dtype=np.int32); flags = np.ascontiguousarray(flags, dtype=np.uint8)
    mp_desc = np.ascontiguousarray(mp_desc, dtype=np.uint8); desc_cur = np.ascontiguousarray(desc_cur, dtype=np.uint8)
    kps_cur = np.ascontiguousarray(kps_cur)
    assert kps_cur.dtype.itemsize == 28
    sf = np.ascontiguousarray(scale_factors, dtype=np.float32)
    ur = None if cur_u_right is None else np.ascontiguousarray(cur_u_right, dtype=np.float32)
    tk = None if cur_taken is None else np.ascontiguousarray(cur_taken, dtype=np.uint8)
    n_mp, n_cur = len(track), len(kps_cur)
    out = np.full(max(n_cur, 1), -1, dtype=np.int32)
    fn.restype = C.c_int
    fn.argtypes = [_f32p, _i32p, _u8p, _u8p, C.c_int, C.c_void_p, _u8p, _f32p, _u8p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                   _f32p, C.c_int, C.c_float, C.c_int, C.c_float, C.c_float, _i32p]
    n = fn(_ptr(track, _f32p), _ptr(level, _i32p), _ptr(flags), _ptr(mp_desc), n_mp, kps_cur.ctypes.data_as(C.c_void_p), _ptr(desc_cur),
           None if ur is None else _ptr(ur, _f32p), None if tk is None else _ptr(tk), n_cur, *[float(b) for b in bounds], _ptr(sf, _f32p),
           len(sf), float(th), 1 if far_points else 0, float(th_far_points), float(nnratio), _ptr(out, _i32p))
    return int(n), out[:n_cur]


def search_by_projection_map(*a, **k):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints) (ORBmatcher.cc:43-210), Nleft == -1 frames ->
    (nmatches, match_cur[n_cur]: index of the map point the feature holds, or -1).  track: n x 5 (mTrackProjX, mTrackProjY, mTrackProjXR,
    mTrackViewCos, mTrackDepth); flags bit 0 = mbTrackInView, bit 1 = Observations() > 0, bit 2 = isBad()."""
    return _search_by_projection_map(port_lib().orbo_search_by_projection_map, *a, **k)


def ref_search_by_projection_map(*a, **k):
    """The same through the reference's own body (librefslices.so)."""
    return _search_by_projection_map(_slices().refslice_search_by_projection_map, *a, **k)


def stage_gray(img, rgb=True, new_size=None):
    """System.cc:420-425 (cv::resize on the interleaved colour image) then Tracking.cc:1569-1582 (cvtColor -> GRAY);
    OpenCV 8-bit RGB2Gray = (R*9798 + G*19235 + B*3735 + 2^14) >> 15."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    planes = [img] if img.ndim == 2 else [np.ascontiguousarray(img[..., c]) for c in range(img.shape[2])]
    if new_size is not None and tuple(new_size) != (img.shape[1], img.shape[0]):
        planes = [resize_linear(p, new_size[0], new_size[1]) for p in planes]
    if len(planes) == 1:
        return planes[0]
    r, g, b = (planes[0], planes[1], planes[2]) if rgb else (planes[2], planes[1], planes[0])
    v = r.astype(np.int64) * 9798 + g.astype(np.int64) * 19235 + b.astype(np.int64) * 3735 + (1 << 14)
    return (v >> 15).astype(np.uint8)


def extract_many(imgs, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, lap=(0, 1000), threads=0,
                 mode="strict"):
    imgs = np.ascontiguousarray(imgs, dtype=np.uint8)
    n, h, w = imgs.shape
    return port_lib().orbo_extract_many(nfeatures, scale_factor, nlevels, ini_th, min_th, 0 if mode == "strict" else 1,
                                        _ptr(imgs), n, w, h, lap[0], lap[1], threads)


def resize_linear(src, dw, dh):
    src = np.ascontiguousarray(src, dtype=np.uint8)
    dst = np.zeros((dh, dw), dtype=np.uint8)
    port_lib().minicv_resize_linear(_ptr(src), src.shape[1], src.shape[0], _ptr(dst), dw, dh)
    return dst


def border101(src, b=19):
    src = np.ascontiguousarray(src, dtype=np.uint8)
    dst = np.zeros((src.shape[0] + 2 * b, src.shape[1] + 2 * b), dtype=np.uint8)
    port_lib().minicv_border101(_ptr(src), src.shape[1], src.shape[0], _ptr(dst), b)
    return dst


def fast(src, threshold, nms=True):
    src = np.ascontiguousarray(src, dtype=np.uint8)
    cap = src.size // 2 + 16
    out = np.zeros((cap, 3), dtype=np.int32)
    n = port_lib().minicv_fast(_ptr(src), src.shape[1], src.shape[0], src.shape[1], threshold, int(nms), _ptr(out, _i32p), cap)
    return out[:n].copy()


def gauss7(src):
    src = np.ascontiguousarray(src, dtype=np.uint8)
    dst = np.zeros_like(src)
    port_lib().minicv_gauss7(_ptr(src), src.shape[1], src.shape[0], _ptr(dst))
    return dst


def fast_atan2(y, x):
    y = np.ascontiguousarray(y, dtype=np.float32); x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.zeros_like(y)
    port_lib().minicv_fast_atan2(_ptr(y, _f32p), _ptr(x, _f32p), _ptr(out, _f32p), len(y))
    return out


# ---- Frame::ComputeBoW (DBoW2 transform) ------------------------------------------------------------------------------

def bow_ref_path():
    return os.path.join(HERE, "_ref", "libbowref.so")


def have_bow_ref():
    return os.path.exists(bow_ref_path())


def write_voc_text(path, parent, node_desc, weight, k, L):
    """ORBvoc.txt format read by the reference's loadFromTextFile (TemplatedVocabulary.h:1340-1420): header "k L scoring
    weighting", then one line per node 1..n-1: parent, isLeaf, 32 descriptor bytes, weight.  NO trailing newline: the
    reference's `while(!f.eof())` loop would turn an empty last line into an extra child of the root."""
    parent = np.asarray(parent)
    n = len(parent)
    has_child = np.zeros(n, bool)
    has_child[parent[1:]] = True
    lines = [f"{k} {L} 0 0"]
    for i in range(1, n):
        d = " ".join(str(int(b)) for b in node_desc[i])
        lines.append(f"{int(parent[i])} {0 if has_child[i] else 1} {d} {float(weight[i])!r}")
    with open(path, "w") as f:
        f.write("\n".join(lines))


def _bow_outputs(n):
    n = max(int(n), 1)
    return (np.zeros(n, np.int32), np.zeros(n, np.float64), np.zeros(1, np.int32), np.zeros(n, np.int32),
            np.zeros(n + 1, np.int32), np.zeros(n, np.int32), np.zeros(1, np.int32))


def _bow_pack(o):
    wi, wv, nw, ni, nb, fi, nn = o
    nw, nn = int(nw[0]), int(nn[0])
    return dict(word_ids=wi[:nw].copy(), word_vals=wv[:nw].copy(), node_ids=ni[:nn].copy(), node_begin=nb[:nn + 1].copy(),
                feat_idx=fi[:int(nb[nn])].copy())


class RefVocabulary:
    """The reference's own DBoW2 vocabulary (oracle/_ref/libbowref.so), loaded from a text file."""

    def __init__(self, path):
        self.lib = C.CDLL(bow_ref_path())
        self.lib.bowref_load_text.restype = C.c_void_p
        self.lib.bowref_load_text.argtypes = [C.c_char_p]
        self.lib.bowref_destroy.argtypes = [C.c_void_p]
        self.lib.bowref_words.argtypes = [C.c_void_p]
        self.lib.bowref_transform.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_int] + [C.c_void_p] * 7
        self.h = self.lib.bowref_load_text(path.encode())
        if not self.h:
            raise RuntimeError("reference loadFromTextFile failed")

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.bowref_destroy(self.h)
            self.h = None

    def words(self):
        return self.lib.bowref_words(self.h)

    def transform(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        o = _bow_outputs(len(desc))
        self.lib.bowref_transform(self.h, _ptr(desc), len(desc), levelsup, *[a.ctypes.data for a in o])
        return _bow_pack(o)


def bow_transform(parent, node_desc, weight, L, desc, levelsup=4):
    """CPU restatement (oracle/bow_oracle.cpp) on the flat vocabulary arrays."""
    lib = port_lib()
    parent = np.ascontiguousarray(parent, np.int32)
    node_desc = np.ascontiguousarray(node_desc, np.uint8).reshape(-1, 32)
    weight = np.ascontiguousarray(weight, np.float64)
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    o = _bow_outputs(len(desc))
    lib.orbo_bow_transform.argtypes = [C.c_int, C.c_void_p, _u8p, C.c_void_p, C.c_int, _u8p, C.c_int, C.c_int] + [C.c_void_p] * 7
    rc = lib.orbo_bow_transform(len(parent), parent.ctypes.data, _ptr(node_desc), weight.ctypes.data, L, _ptr(desc), len(desc),
                                levelsup, *[a.ctypes.data for a in o])
    if rc != 0:
        raise ValueError("malformed vocabulary tree")
    return _bow_pack(o)


class PortVocabulary:
    """oracle/bow_oracle.cpp with the child lists kept between calls (what bench.py's cpu_baseline leg times)."""

    def __init__(self, parent, node_desc, weight, L):
        lib = port_lib()
        parent = np.ascontiguousarray(parent, np.int32)
        node_desc = np.ascontiguousarray(node_desc, np.uint8).reshape(-1, 32)
        weight = np.ascontiguousarray(weight, np.float64)
        lib.orbo_bow_create.restype = C.c_void_p
        lib.orbo_bow_create.argtypes = [C.c_int, C.c_void_p, _u8p, C.c_void_p, C.c_int]
        lib.orbo_bow_destroy.argtypes = [C.c_void_p]
        lib.orbo_bow_transform_h.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_int] + [C.c_void_p] * 7
        self.lib = lib
        self.h = lib.orbo_bow_create(len(parent), parent.ctypes.data, _ptr(node_desc), weight.ctypes.data, L)
        if not self.h:
            raise ValueError("malformed vocabulary tree")

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.orbo_bow_destroy(self.h)
            self.h = None

    def transform(self, desc, levelsup=4):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        o = _bow_outputs(len(desc))
        rc = self.lib.orbo_bow_transform_h(self.h, _ptr(desc), len(desc), levelsup, *[a.ctypes.data for a in o])
        if rc != 0:
            raise ValueError("malformed vocabulary tree")
        return _bow_pack(o)


# ---- the reference's own Frame.cc / ORBmatcher.cc function bodies (oracle/_ref/librefslices.so) ------------------------

def refslices_path():
    return os.path.join(HERE, "_ref", "librefslices.so")


def have_refslices():
    return os.path.exists(refslices_path())


_refslices = None


def _slices():
    global _refslices
    if _refslices is None:
        lib = C.CDLL(refslices_path())
        lib.refslice_descriptor_distance.argtypes = [_u8p, _u8p]
        lib.refslice_compute_stereo_matches.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_void_p, _u8p, C.c_int, _f32p, _f32p, C.c_int,
                                                        C.c_void_p, C.c_void_p, _i32p, _i32p, C.c_float, C.c_float, _f32p, _f32p]
        lib.refslice_grid_top2.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, _u8p,
                                           C.c_int, _i32p, _i32p, _i32p, _i32p, _i32p]
        lib.refslice_undistort_keypoints.argtypes = [C.c_void_p, C.c_int, _f32p, _f32p, C.c_int, _f32p, C.c_void_p]
        lib.refslice_image_bounds.argtypes = [C.c_int, C.c_int, _f32p, _f32p, C.c_int, _f32p, _f32p]
        lib.refslice_accept.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        lib.refslice_accept_grid.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, _u8p]
        _refslices = lib
    return _refslices


def ref_accept_grid(site, ratio, level1=0, level2=0, results=2):
    """The reference's own accept condition at ORBmatcher.cc:<site> (Frame.cc:1151 for site 1151), compiled from the sliced source
    lines (librefslices.so), over every (d1, d2) in [0, 256]^2 -> bool array [257, 257]."""
    out = np.zeros((257, 257), dtype=np.uint8)
    _slices().refslice_accept_grid(int(site), float(ratio), int(level1), int(level2), int(results), _ptr(out))
    assert out.max() <= 1, "unknown site"
    return out.astype(bool)


def ref_thresholds():
    """(TH_LOW, TH_HIGH) as the reference's ORBmatcher.cc:35-37 defines them."""
    return _slices().refslice_th_low(), _slices().refslice_th_high()


def ref_descriptor_distance(a, b):
    """ORBmatcher::DescriptorDistance (ORBmatcher.cc:2058-2074), the reference's own body."""
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    return _slices().refslice_descriptor_distance(_ptr(a), _ptr(b))


def ref_compute_stereo_matches(kl, dl, kr, dr, scale_factors, inv_scale_factors, pyr_l, pyr_r, bf, b):
    """Frame::ComputeStereoMatches (Frame.cc:811-981), the reference's own body; arguments as compute_stereo_matches."""
    kl = np.ascontiguousarray(kl); kr = np.ascontiguousarray(kr)
    dl = np.ascontiguousarray(dl, dtype=np.uint8); dr = np.ascontiguousarray(dr, dtype=np.uint8)
    sf = np.ascontiguousarray(scale_factors, dtype=np.float32); isf = np.ascontiguousarray(inv_scale_factors, dtype=np.float32)
    nl = len(sf)
    pl = [np.ascontiguousarray(p, dtype=np.uint8) for p in pyr_l]
    pr = [np.ascontiguousarray(p, dtype=np.uint8) for p in pyr_r]
    ptr_l = (C.c_void_p * nl)(*[p.ctypes.data for p in pl])
    ptr_r = (C.c_void_p * nl)(*[p.ctypes.data for p in pr])
    w = np.array([p.shape[1] - 38 for p in pl], dtype=np.int32); h = np.array([p.shape[0] - 38 for p in pl], dtype=np.int32)
    n = len(kl)
    ur = np.zeros(max(n, 1), dtype=np.float32); dp = np.zeros(max(n, 1), dtype=np.float32)
    _slices().refslice_compute_stereo_matches(kl.ctypes.data_as(C.c_void_p), _ptr(dl), n, kr.ctypes.data_as(C.c_void_p), _ptr(dr), len(kr),
                                              _ptr(sf, _f32p), _ptr(isf, _f32p), nl, ptr_l, ptr_r, _ptr(w, _i32p), _ptr(h, _i32p),
                                              float(bf), float(b), _ptr(ur, _f32p), _ptr(dp, _f32p))
    return ur[:n], dp[:n]


def ref_grid_top2(kps_t, desc_t, bounds, queries, desc_q):
    """AssignFeaturesToGrid + GetFeaturesInArea + PosInGrid (Frame.cc:385-416, 657-735, the reference's own bodies) and the
    strict-< best/second loop over what they return -> idx1, d1, idx2, d2, candidates per query.  queries: the 24-byte
    orbx_grid_query records (ur is ignored: monocular frame)."""
    kps_t = np.ascontiguousarray(kps_t); queries = np.ascontiguousarray(queries)
    desc_t = np.ascontiguousarray(desc_t, dtype=np.uint8); desc_q = np.ascontiguousarray(desc_q, dtype=np.uint8)
    nq = len(queries)
    q20 = np.zeros((nq, 20), np.uint8)
    if nq:
        q20[:] = queries.view(np.uint8).reshape(nq, 24)[:, :20]
    out = [np.zeros(max(nq, 1), dtype=np.int32) for _ in range(5)]
    _slices().refslice_grid_top2(kps_t.ctypes.data_as(C.c_void_p), _ptr(desc_t), len(kps_t), *[float(b) for b in bounds],
                                 q20.ctypes.data, _ptr(desc_q), nq, *[_ptr(o, _i32p) for o in out])
    return [o[:nq] for o in out]


def _search_by_bow(fn, desc_kf, valid_kf, angle_kf, fv_kf, desc_f, angle_f, fv_f, nnratio, check_orientation):
    desc_kf = np.ascontiguousarray(desc_kf, np.uint8).reshape(-1, 32); desc_f = np.ascontiguousarray(desc_f, np.uint8).reshape(-1, 32)
    valid_kf = np.ascontiguousarray(valid_kf, np.uint8)
    angle_kf = np.ascontiguousarray(angle_kf, np.float32); angle_f = np.ascontiguousarray(angle_f, np.float32)
    arrs = []
    for fv in (fv_kf, fv_f):
        arrs.append([np.ascontiguousarray(fv[k], np.int32) for k in ("node_ids", "node_begin", "feat_idx")])
    out = np.full(max(len(desc_f), 1), -1, np.int32)
    fn.restype = C.c_int
    fn.argtypes = [_u8p, C.c_int, _u8p, _f32p, _i32p, _i32p, _i32p, C.c_int, _u8p, C.c_int, _f32p, _i32p, _i32p, _i32p, C.c_int,
                   C.c_float, C.c_int, _i32p]
    n = fn(_ptr(desc_kf), len(desc_kf), _ptr(valid_kf), _ptr(angle_kf, _f32p), *[_ptr(a, _i32p) for a in arrs[0]], len(arrs[0][0]),
           _ptr(desc_f), len(desc_f), _ptr(angle_f, _f32p), *[_ptr(a, _i32p) for a in arrs[1]], len(arrs[1][0]),
           float(nnratio), int(bool(check_orientation)), _ptr(out, _i32p))
    return n, out[:len(desc_f)]


def _search_by_bow_kf(fn, desc1, valid1, angle1, fv1, desc2, valid2, angle2, fv2, nnratio, check_orientation):
    desc1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32); desc2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
    valid1 = np.ascontiguousarray(valid1, np.uint8); valid2 = np.ascontiguousarray(valid2, np.uint8)
    angle1 = np.ascontiguousarray(angle1, np.float32); angle2 = np.ascontiguousarray(angle2, np.float32)
    arrs = []
    for fv in (fv1, fv2):
        arrs.append([np.ascontiguousarray(fv[k], np.int32) for k in ("node_ids", "node_begin", "feat_idx")])
    out = np.full(max(len(desc1), 1), -1, np.int32)
    fn.restype = C.c_int
    fn.argtypes = [_u8p, C.c_int, _u8p, _f32p, _i32p, _i32p, _i32p, C.c_int, _u8p, C.c_int, _u8p, _f32p, _i32p, _i32p, _i32p, C.c_int,
                   C.c_float, C.c_int, _i32p]
    n = fn(_ptr(desc1), len(desc1), _ptr(valid1), _ptr(angle1, _f32p), *[_ptr(a, _i32p) for a in arrs[0]], len(arrs[0][0]),
           _ptr(desc2), len(desc2), _ptr(valid2), _ptr(angle2, _f32p), *[_ptr(a, _i32p) for a in arrs[1]], len(arrs[1][0]),
           float(nnratio), int(bool(check_orientation)), _ptr(out, _i32p))
    return n, out[:len(desc1)]


def search_by_bow_kf(desc1, valid1, angle1, fv1, desc2, valid2, angle2, fv2, nnratio=0.75, check_orientation=True):
    """ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, ...) restatement (ORBmatcher.cc:765-905) -> (nmatches, match12)."""
    return _search_by_bow_kf(port_lib().orbo_search_by_bow_kf, desc1, valid1, angle1, fv1, desc2, valid2, angle2, fv2, nnratio,
                             check_orientation)


def ref_search_by_bow_kf(desc1, valid1, angle1, fv1, desc2, valid2, angle2, fv2, nnratio=0.75, check_orientation=True):
    """The same through the reference's own body (librefslices.so)."""
    return _search_by_bow_kf(_slices().refslice_search_by_bow_kf, desc1, valid1, angle1, fv1, desc2, valid2, angle2, fv2, nnratio,
                             check_orientation)


def search_by_bow(desc_kf, valid_kf, angle_kf, fv_kf, desc_f, angle_f, fv_f, nnratio=0.7, check_orientation=True):
    """ORBmatcher::SearchByBoW(KF, F) restatement (oracle/match_oracle.cpp) -> (nmatches, match_f).  fv_*: dicts with the
    FeatureVector CSR arrays of a BoW transform (node_ids, node_begin, feat_idx)."""
    return _search_by_bow(port_lib().orbo_search_by_bow, desc_kf, valid_kf, angle_kf, fv_kf, desc_f, angle_f, fv_f, nnratio, check_orientation)


def ref_search_by_bow(desc_kf, valid_kf, angle_kf, fv_kf, desc_f, angle_f, fv_f, nnratio=0.7, check_orientation=True):
    """The same through the reference's own function body (ORBmatcher.cc:223-425, oracle/_ref/librefslices.so)."""
    return _search_by_bow(_slices().refslice_search_by_bow, desc_kf, valid_kf, angle_kf, fv_kf, desc_f, angle_f, fv_f, nnratio, check_orientation)
