#!/usr/bin/env python3
"""Generate the committed golden fixtures under tests/golden/ (run in the BUILD container only).

  prim_*.npz   outputs of cv2 (OpenCV 4.13.0, the declared oracle version) for the five
               un-vendored primitives on small inputs -> pins oracle/minicv
  prim_undistort.npz  cv2.undistortPoints on the pinhole cases of synth.UNDISTORT_CASES -> pins minicv's undistortPoints
  match.npz    cv2.BFMatcher(NORM_HAMMING).knnMatch(k=2) on descriptor sets with planted ties
               -> pins the matcher oracle's (distance, index) order (reference: ORB_SLAM3/src/Frame.cc:1144)
  bow.npz      BowVector / FeatureVector from the reference's own DBoW2 (oracle/_ref/libbowref.so) on synthetic
               vocabulary trees -> pins the Frame::ComputeBoW oracle and is the GPU parity fixture
  e2e_*.npz    keypoints / descriptors / pyramid digests from the reference's own ORBextractor.cc compiled
               unchanged on the minicv shim (oracle/_ref) on synthetic frames -> pins the port oracle and
               is what the GPU parity tests compare against on the GPU box, where /root/reference is absent.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orb_slam3_ros2_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def level_sizes(w, h, nlevels=8, sf=1.2):
    s = np.float32(1.0)
    out = []
    for l in range(nlevels):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
        s = np.float32(np.float64(s) * np.float64(np.float32(sf)))
    return out


def gen_primitives():
    import cv2
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    rng = np.random.default_rng(1234)
    imgs = {
        "rand": rng.integers(0, 256, (120, 160), dtype=np.uint8),
        "smooth": cv2.GaussianBlur(rng.integers(0, 256, (97, 131), dtype=np.uint8), (9, 9), 3),
        "synth": synth.frame(3, 200, 150),
        "odd": rng.integers(0, 256, (61, 77), dtype=np.uint8),
    }
    # resize chains exactly as ComputePyramid does (sizes from the original, source = previous level)
    rz = {}
    for name, im in imgs.items():
        prev = im
        for l, (w, h) in enumerate(level_sizes(im.shape[1], im.shape[0])[1:5], start=1):
            cur = cv2.resize(prev, (w, h), interpolation=cv2.INTER_LINEAR)
            rz[f"{name}_L{l}"] = cur
            prev = cur
        rz[f"{name}_src"] = im
    np.savez_compressed(os.path.join(OUT, "prim_resize.npz"), **rz)
    bd = {}
    for name, im in imgs.items():
        bd[f"{name}_src"] = im
        bd[f"{name}_b19"] = cv2.copyMakeBorder(im, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    np.savez_compressed(os.path.join(OUT, "prim_border.npz"), **bd)
    fs = {}
    fast_imgs = dict(imgs)
    fast_imgs["cell"] = synth.frame(1, 42, 44)
    fast_imgs["cell2"] = cv2.GaussianBlur(rng.integers(0, 256, (41, 47), dtype=np.uint8), (3, 3), 1)
    fast_imgs["tiny7"] = rng.integers(0, 256, (7, 7), dtype=np.uint8)
    fast_imgs["tiny8"] = rng.integers(0, 256, (9, 8), dtype=np.uint8)
    fast_imgs["flat"] = np.full((30, 30), 100, dtype=np.uint8)
    for name, im in fast_imgs.items():
        fs[f"{name}_src"] = im
        for th in (20, 7, 0, 60):
            det = cv2.FastFeatureDetector_create(th, True)
            kps = det.detect(im)
            arr = np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in kps], dtype=np.int32).reshape(-1, 3)
            assert all(k.size == 7 and k.angle == -1 and k.octave == 0 for k in kps)
            fs[f"{name}_t{th}"] = arr
    np.savez_compressed(os.path.join(OUT, "prim_fast.npz"), **fs)
    bl = {}
    for name, im in imgs.items():
        bl[f"{name}_src"] = im
        bl[f"{name}_g7"] = cv2.GaussianBlur(im, (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101)
    small = rng.integers(0, 256, (5, 9), dtype=np.uint8)
    bl["small_src"] = small
    bl["small_g7"] = cv2.GaussianBlur(small, (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101)
    np.savez_compressed(os.path.join(OUT, "prim_blur.npz"), **bl)
    # fastAtan2 on integer-valued moments (what IC_Angle feeds it) plus generic floats and specials
    y = np.concatenate([rng.integers(-2700000, 2700000, 3000).astype(np.float32),
                        rng.normal(0, 100, 1000).astype(np.float32),
                        np.array([0, 0, -5, 5, 0, 1, -1, 1e-30, 3], dtype=np.float32)])
    x = np.concatenate([rng.integers(-2700000, 2700000, 3000).astype(np.float32),
                        rng.normal(0, 100, 1000).astype(np.float32),
                        np.array([0, 5, 0, 0, -3, 1, -1, 1e-30, -3], dtype=np.float32)])
    ang = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], dtype=np.float32)
    np.savez_compressed(os.path.join(OUT, "prim_atan2.npz"), y=y, x=x, angle=ang)
    # knnMatch k=2 with ties / duplicates / distance 0 and 256
    q = rng.integers(0, 256, (64, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    t[10] = q[0]; t[200] = q[0]            # exact duplicates (d=0 twice) -> lowest index first
    t[17] = ~q[1]                          # distance 256
    t[50] = t[40]                          # duplicate train rows -> equal distances for every query
    q[2] = t[40]
    bf = cv2.BFMatcher(cv2.NORM_HAMMING)
    m = bf.knnMatch(q, t, 2)
    idx = np.array([[mm[0].trainIdx, mm[1].trainIdx] for mm in m], dtype=np.int32)
    dist = np.array([[mm[0].distance, mm[1].distance] for mm in m], dtype=np.float32)
    m1 = bf.knnMatch(q, t[:1], 2)
    assert all(len(mm) == 1 for mm in m1)
    np.savez_compressed(os.path.join(OUT, "match.npz"), q=q, t=t, idx=idx, dist=dist)
    print("primitives + match goldens written (cv2", cv2.__version__ + ")")


STAGE_CASES = [
    # name, source (w, h), channels, rgb (mbRGB), new size or None
    ("rgb_same", (97, 61), 3, True, None),
    ("bgr_same", (97, 61), 3, False, None),
    ("rgba_same", (64, 48), 4, True, None),
    ("bgra_half", (128, 96), 4, False, (64, 48)),
    ("rgb_down", (160, 120), 3, True, (94, 60)),
    ("bgr_up", (80, 60), 3, False, (101, 77)),
    ("gray_down", (131, 97), 1, True, (100, 64)),
    ("rgb_half", (120, 90), 3, True, (60, 45)),
]


def gen_stage():
    """cv2.resize (on the interleaved colour image) then cv2.cvtColor(-> GRAY): System.cc:420-425 + Tracking.cc:1569-1582."""
    import cv2
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    rng = np.random.default_rng(77)
    rec = {}
    codes = {(3, True): cv2.COLOR_RGB2GRAY, (3, False): cv2.COLOR_BGR2GRAY, (4, True): cv2.COLOR_RGBA2GRAY,
             (4, False): cv2.COLOR_BGRA2GRAY}
    for name, (w, h), ch, rgb, new in STAGE_CASES:
        shape = (h, w) if ch == 1 else (h, w, ch)
        im = rng.integers(0, 256, shape, dtype=np.uint8)
        if name.endswith("half"):  # smooth content too
            im = cv2.GaussianBlur(im, (5, 5), 1.5)
        out = im if new is None else cv2.resize(im, new)
        if ch != 1:
            out = cv2.cvtColor(out, codes[(ch, rgb)])
        rec[name + "_src"] = im
        rec[name + "_gray"] = out
        rec[name + "_meta"] = np.array([ch, int(rgb), new[0] if new else 0, new[1] if new else 0])
    # every (r, g, b) on a lattice that includes the extremes: pins the grey weights and the rounding
    vals = np.array(sorted(set(list(range(0, 256, 9)) + [1, 127, 128, 254, 255])), dtype=np.uint8)
    R, G, B = np.meshgrid(vals, vals, vals, indexing="ij")
    cube = np.stack([R.ravel(), G.ravel(), B.ravel()], 1).reshape(1, -1, 3)
    rec["cube_src"] = cube
    rec["cube_gray"] = cv2.cvtColor(cube, cv2.COLOR_RGB2GRAY)
    rec["cube_meta"] = np.array([3, 1, 0, 0])
    np.savez_compressed(os.path.join(OUT, "stage.npz"), **rec)
    print("staging goldens written (cv2", cv2.__version__ + ")")


E2E_CASES = [
    # name, w, h, frame k, nfeatures, lapping area, right
    ("euroc752", 752, 480, 0, 1000, (0, 1000), False),
    ("d435i640", 640, 480, 5, 1000, (0, 1000), False),
    ("kittiL", 1241, 376, 2, 2000, (0, 0), False),
    ("kittiR", 1241, 376, 2, 2000, (0, 0), True),
    ("hd1080", 1920, 1080, 1, 4000, (0, 1000), False),
    ("milkv360", 640, 360, 9, 1250, (300, 600), False),
    ("init5x", 752, 480, 4, 5000, (0, 1000), False),
    ("small320", 320, 240, 11, 500, (0, 1000), False),
]


def gen_e2e():
    from oracle import pyoracle as po
    po.build()
    assert po.have_ref("strict") and po.have_ref("native"), "oracle/_ref missing: run where /root/reference exists"
    for name, w, h, k, nf, lap, right in E2E_CASES:
        im = synth.frame(k, w, h, 42, right)
        rec = dict(w=w, h=h, k=k, nfeatures=nf, lap=np.array(lap), right=right,
                   img_sha1=hashlib.sha1(im.tobytes()).hexdigest())
        for mode in ("strict", "native"):
            e = po.RefExtractor(nfeatures=nf, mode=mode)
            mono, kp, desc = e.extract(im, lap)
            rec[f"{mode}_mono"] = mono
            rec[f"{mode}_kp"] = kp
            rec[f"{mode}_desc"] = desc
            if mode == "strict":
                rec["pyr_sha1"] = np.array([hashlib.sha1(e.pyramid(l).tobytes()).hexdigest() for l in range(8)])
                rec["pyr_shape"] = np.array([e.pyramid(l).shape for l in range(8)])
        if np.array_equal(rec["strict_desc"], rec["native_desc"]) and np.array_equal(rec["strict_kp"], rec["native_kp"]):
            rec["native_same"] = True
            del rec["native_kp"], rec["native_desc"]
        np.savez_compressed(os.path.join(OUT, f"e2e_{name}.npz"), **rec)
        print(name, "n =", len(rec["strict_kp"]), "mono =", rec["strict_mono"], "native_same =", rec.get("native_same", False))


def gen_undistort():
    """cv2.undistortPoints(pts, K, D, None, K) as Frame::UndistortKeyPoints / ComputeImageBounds call it (Frame.cc:766, :795)."""
    import cv2
    rec = {"cv2_version": cv2.__version__}
    for name, *_ in synth.UNDISTORT_CASES:
        w, h, K, dist, pts = synth.undistort_case(name)
        out = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, dist.reshape(-1, 1), None, K).reshape(-1, 2)
        rec[f"{name}__pts"] = pts
        rec[f"{name}__out"] = out.astype(np.float32)
        print(name, "max shift", float(np.abs(out - pts).max()))
    np.savez_compressed(os.path.join(OUT, "prim_undistort.npz"), **rec)


def gen_bow():
    """Frame::ComputeBoW: outputs of the reference's own DBoW2 (oracle/_ref/libbowref.so) on synthetic vocabularies."""
    import tempfile
    from oracle import pyoracle as po
    po.build()
    assert po.have_bow_ref(), "oracle/_ref/libbowref.so missing: run where /root/reference exists"
    rec = {}
    for name, *_ in synth.BOW_CASES:
        k, L, levelsup, parent, node_desc, weight, desc = synth.bow_case(name)
        with tempfile.TemporaryDirectory() as t:
            path = os.path.join(t, "voc.txt")
            po.write_voc_text(path, parent, node_desc, weight, k, L)
            voc = po.RefVocabulary(path)
            out = voc.transform(desc, levelsup)
            rec[f"{name}__words"] = voc.words()
        rec[f"{name}__voc_sha1"] = hashlib.sha1(parent.tobytes() + node_desc.tobytes() + weight.tobytes() + desc.tobytes()).hexdigest()
        for key, val in out.items():
            rec[f"{name}__{key}"] = val
        print(name, "nodes", len(parent), "words hit", len(out["word_ids"]), "feature-vector nodes", len(out["node_ids"]))
    # the descriptors of a real extraction (the 752x480 fixture) through a 10^4-word tree, levelsup 2
    e2e = np.load(os.path.join(OUT, "e2e_euroc752.npz"))
    k, L, levelsup, parent, node_desc, weight, _ = synth.bow_case("k10L4_levelsup2")
    with tempfile.TemporaryDirectory() as t:
        path = os.path.join(t, "voc.txt")
        po.write_voc_text(path, parent, node_desc, weight, k, L)
        out = po.RefVocabulary(path).transform(e2e["strict_desc"], levelsup)
    for key, val in out.items():
        rec[f"euroc752__{key}"] = val
    np.savez_compressed(os.path.join(OUT, "bow.npz"), **rec)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    what = sys.argv[1:] or ["prim", "stage", "e2e", "bow", "undistort"]
    if "prim" in what:
        gen_primitives()
    if "stage" in what:
        gen_stage()
    if "e2e" in what:
        gen_e2e()
    if "bow" in what:
        gen_bow()
    if "undistort" in what:
        gen_undistort()
