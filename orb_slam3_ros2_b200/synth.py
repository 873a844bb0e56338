"""Deterministic synthetic inputs, integer-only so that numpy, the C oracle
and the CUDA generator (csrc/synth.cu) agree bit-for-bit.

There is no dataset offline (EuRoC/KITTI/RealSense streams are unavailable), so BASELINE.json's
configs are measured on these procedurally textured frames and descriptor sets.
"""
import numpy as np

K1 = np.uint64(0x9E3779B97F4A7C15)
K2 = np.uint64(0xC2B2AE3D27D4EB4F)
K3 = np.uint64(0x165667B19E3779F9)
OCTAVES = ((4, 10), (8, 25), (16, 40), (32, 50), (64, 60))  # (lattice cell, amplitude)
ORIGIN = 1024


def _mix(z):
    z = z.astype(np.uint64)
    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def hash3(seed, a, b, c):
    """splitmix64 finaliser of seed ^ a*K1 ^ b*K2 ^ c*K3 (uint64 wrap-around)."""
    with np.errstate(over="ignore"):
        a = np.asarray(a).astype(np.uint64) * K1
        b = np.asarray(b).astype(np.uint64) * K2
        c = np.asarray(c).astype(np.uint64) * K3
    return _mix(np.uint64(seed) ^ a ^ b ^ c)


def _tdiv(n, d):
    """C-style integer division (truncation toward zero)."""
    q = np.abs(n) // d
    return np.where(n < 0, -q, q)


def canvas(X, Y, seed):
    """I(X, Y; seed) for integer canvas coordinates X, Y >= 64 (arrays broadcast)."""
    X = np.asarray(X, dtype=np.int64)
    Y = np.asarray(Y, dtype=np.int64)
    X, Y = np.broadcast_arrays(X, Y)
    val = np.full(X.shape, 128, dtype=np.int64)
    for o, (c, amp) in enumerate(OCTAVES):
        i, fx = X // c, X % c
        j, fy = Y // c, Y % c

        def g(ii, jj):
            return (hash3(seed, o, ii, jj) & np.uint64(2047)).astype(np.int64) - 1024

        num = (g(i, j) * (c - fx) * (c - fy) + g(i + 1, j) * fx * (c - fy)
               + g(i, j + 1) * (c - fx) * fy + g(i + 1, j + 1) * fx * fy)
        val += _tdiv(amp * num, 1024 * c * c)
    ti, tj = X // 64, Y // 64
    for dti in (-1, 0):
        for dtj in (-1, 0):
            for r in range(3):
                hv = hash3(seed, 100 + r, ti + dti, tj + dtj)
                ox = (hv & np.uint64(63)).astype(np.int64)
                oy = ((hv >> np.uint64(6)) & np.uint64(63)).astype(np.int64)
                w = 4 + ((hv >> np.uint64(12)) % np.uint64(36)).astype(np.int64)
                h = 4 + ((hv >> np.uint64(20)) % np.uint64(36)).astype(np.int64)
                delta = ((hv >> np.uint64(28)) % np.uint64(160)).astype(np.int64) - 80
                filled = ((hv >> np.uint64(40)) % np.uint64(10)) < np.uint64(7)
                x0 = (ti + dti) * 64 + ox
                y0 = (tj + dtj) * 64 + oy
                inside = (X >= x0) & (X < x0 + w) & (Y >= y0) & (Y < y0 + h)
                edge = (X == x0) | (X == x0 + w - 1) | (Y == y0) | (Y == y0 + h - 1)
                val += np.where(inside & (filled | edge), delta, 0)
    return np.clip(val, 0, 255).astype(np.uint8)


def frame(k, width, height, seed=42, right=False):
    """Frame k of the synthetic stream: crop of the canvas at offset (3k, k); the stereo right
    image is the same crop shifted by +16 px in X (so uR = uL - 16)."""
    x = np.arange(width, dtype=np.int64)[None, :] + ORIGIN + 3 * k + (16 if right else 0)
    y = np.arange(height, dtype=np.int64)[:, None] + ORIGIN + k
    return np.ascontiguousarray(canvas(x, y, seed))


def frames(k0, n, width, height, seed=42, right=False):
    return np.stack([frame(k0 + i, width, height, seed, right) for i in range(n)])


def descriptor_db(n, seed=7, start=0):
    """n x 32 bytes: row r, 64-bit word j (little endian) = splitmix64 finaliser of hash3(seed, r, j, 0)."""
    r = np.arange(start, start + n, dtype=np.uint64)[:, None]
    j = np.arange(4, dtype=np.uint64)[None, :]
    words = hash3(seed, r, j, 0)
    return np.ascontiguousarray(words).view(np.uint8).reshape(n, 32)


def queries_from_db(nq, n_db, db_seed=7, q_seed=11, flips=20):
    """Query i = DB row (i*4999+17) mod n_db with `flips` distinct pseudo-random bit flips,
    so the true nearest neighbour and its distance (= flips) are known."""
    rows = (np.arange(nq, dtype=np.int64) * 4999 + 17) % n_db
    q = np.empty((nq, 32), dtype=np.uint8)
    for i, r in enumerate(rows):
        d = descriptor_db(1, db_seed, int(r))[0].copy()
        bits = set()
        t = 0
        while len(bits) < flips:
            bits.add(int(hash3(q_seed, i, t, 1)) & 255)
            t += 1
        for b in bits:
            d[b >> 3] ^= np.uint8(1 << (b & 7))
        q[i] = d
    return q, rows


def vocabulary(k=10, L=3, seed=5, ragged=False, stop_every=0, flips=24, leaf_level=None):
    """Synthetic ORB vocabulary tree in the flat form of orbx_voc_create (ORBvoc.txt.tar.gz is a missing blob offline,
    .MISSING_LARGE_BLOBS): node 0 = root, nodes in breadth-first order so parent[i] < i, every child descriptor = its parent's
    with `flips` pseudo-random bits flipped (the descent is then meaningful, like a k-means tree), idf-like weights.
    ragged: inner nodes get between 2 and k children.  stop_every: every n-th word gets weight 0 (a stopped word,
    TemplatedVocabulary.h:1157).  leaf_level: leaves may already appear at this depth (< L) for ~1/7 of the nodes there.
    Returns (parent int32[n], node_desc uint8[n,32], weight float64[n])."""
    rng = np.random.default_rng(seed)
    parent, desc, depth = [-1], [np.zeros(32, np.uint8)], [0]
    frontier = [0]
    for level in range(1, L + 1):
        nxt = []
        for p in frontier:
            if leaf_level is not None and depth[p] >= leaf_level and p % 7 == 3:
                continue  # an early leaf
            nc = int(rng.integers(2, k + 1)) if ragged else k
            for _ in range(nc):
                d = desc[p].copy() if p else rng.integers(0, 256, 32, dtype=np.uint8)
                if p:
                    for b in rng.integers(0, 256, flips):
                        d[b >> 3] ^= np.uint8(1 << (b & 7))
                parent.append(p)
                desc.append(d)
                depth.append(level)
                nxt.append(len(parent) - 1)
        frontier = nxt
    n = len(parent)
    parent = np.asarray(parent, np.int32)
    weight = np.round(rng.uniform(0.5, 9.0, n), 6)  # 6 decimals: survives the text file round trip exactly
    has_child = np.zeros(n, bool)
    has_child[parent[1:]] = True
    if stop_every:
        leaves = np.flatnonzero(~has_child)
        weight[leaves[::stop_every]] = 0.0
    weight[has_child] = 0.0
    return parent, np.stack(desc), weight


def descriptors_near_words(parent, node_desc, n, seed=9, flips=30):
    """n query descriptors: random leaves of the vocabulary with `flips` bits flipped (plus a few exact copies, so that
    distance-0 and tie cases occur)."""
    rng = np.random.default_rng(seed)
    has_child = np.zeros(len(parent), bool)
    has_child[np.asarray(parent)[1:]] = True
    leaves = np.flatnonzero(~has_child)
    out = node_desc[rng.choice(leaves, n)].copy()
    for i in range(n):
        if i % 17 == 0:
            continue
        for b in rng.integers(0, 256, flips):
            out[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return out


# (name, k, L, ragged, stop_every, leaf_level, levelsup, n descriptors): the Frame::ComputeBoW parity cases shared by
# tools/gen_golden.py (reference DBoW2 -> tests/golden/bow.npz) and the CPU / GPU tests
BOW_CASES = [
    ("k10L3", 10, 3, False, 0, None, 1, 1000),
    ("k10L3_ragged_stop", 10, 3, True, 5, None, 2, 1000),
    ("k6L4_ragged_root", 6, 4, True, 3, None, 4, 700),
    ("k10L4_earlyleaf", 10, 4, False, 7, 3, 1, 1500),
    ("k3L2_tiny", 3, 2, False, 0, None, 4, 64),
    ("k10L4_levelsup2", 10, 4, False, 0, None, 2, 2000),
    ("k10L3_one", 10, 3, False, 0, None, 1, 1),
]

# (name, w, h, fx, fy, cx, cy, distortion (k1, k2, p1, p2[, k3])): the Frame::UndistortKeyPoints / ComputeImageBounds parity
# cases (pinhole cameras) shared by tools/gen_golden.py (cv2.undistortPoints -> tests/golden/prim_undistort.npz) and the tests.
# euroc = the reference's config/Monocular-Inertial/EuRoC.yaml:11-19; the others exercise k3, strong barrel / pincushion
# distortion, the icdist < 0 exit and the zero-k1 early return (Frame.cc:749-753).
UNDISTORT_CASES = [
    ("euroc", 752, 480, 458.654, 457.296, 367.215, 248.375, (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05)),
    ("k3", 640, 480, 382.613, 382.613, 320.183, 236.455, (-0.05, 0.04, 0.001, -0.002, 0.01)),
    ("kitti_like", 1241, 376, 718.856, 718.856, 607.1928, 185.2157, (0.1, -0.2, 0.0005, 0.0003)),
    ("strong_barrel", 752, 480, 300.0, 300.0, 376.0, 240.0, (-0.9, 0.1, 0.0, 0.0)),
    ("k1_zero", 640, 480, 382.613, 382.613, 320.183, 236.455, (0.0, 0.04, 0.001, -0.002)),
]


def undistort_case(name, n=4000):
    """-> (w, h, K float32 3x3, dist float32, points float32 n x 2): image corners first (the ComputeImageBounds input,
    Frame.cc:786-792), then seeded points inside the image, a few on its border and a few far outside."""
    for nm, w, h, fx, fy, cx, cy, dist in UNDISTORT_CASES:
        if nm == name:
            rng = np.random.default_rng(abs(hash_name(nm)) % (1 << 32))
            K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
            pts = (rng.random((n, 2)) * [w, h]).astype(np.float32)
            pts[:4] = [[0, 0], [w, 0], [0, h], [w, h]]
            pts[4:8] = [[cx, cy], [w - 1, h - 1], [-50.5, 3 * h], [2.5 * w, -h]]
            return w, h, K, np.array(dist, np.float32), pts
    raise KeyError(name)


def hash_name(s):
    v = 0
    for ch in s.encode():
        v = (v * 131 + ch) & 0x7FFFFFFF
    return v


def bow_case(name):
    """-> (k, L, levelsup, parent, node_desc, weight, descriptors) of a BOW_CASES entry."""
    for nm, k, L, ragged, stop, leaf_level, levelsup, n in BOW_CASES:
        if nm == name:
            parent, node_desc, weight = vocabulary(k, L, seed=5, ragged=ragged, stop_every=stop, leaf_level=leaf_level)
            return k, L, levelsup, parent, node_desc, weight, descriptors_near_words(parent, node_desc, n)
    raise KeyError(name)


def vocabulary_full(k=10, L=6, seed=5, flips=24):
    """A complete k-ary tree of depth L in breadth-first order, vectorised (the ORBvoc.txt shape k = 10, L = 6 has
    1,111,111 nodes / 10^6 words): same construction as vocabulary(), level by level with numpy."""
    rng = np.random.default_rng(seed)
    parents, descs = [np.array([-1], np.int64)], [np.zeros((1, 32), np.uint8)]
    first, count = 0, 1
    for level in range(1, L + 1):
        n = count * k
        par = first + np.repeat(np.arange(count, dtype=np.int64), k)
        if level == 1:
            d = rng.integers(0, 256, (n, 32), dtype=np.uint8)
        else:
            d = np.repeat(descs[-1], k, axis=0)
            pos = rng.integers(0, 256, (n, flips))
            rows = np.repeat(np.arange(n), flips)
            np.bitwise_xor.at(d, (rows, (pos >> 3).ravel()), (1 << (pos & 7)).astype(np.uint8).ravel())
        parents.append(par)
        descs.append(d)
        first, count = first + count, n
    parent = np.concatenate(parents).astype(np.int32)
    node_desc = np.concatenate(descs)
    weight = np.round(rng.uniform(0.5, 9.0, len(parent)), 6)
    weight[:first] = 0.0
    return parent, node_desc, weight
