import sys, numpy as np
sys.path.insert(0, '.')
from orb_slam3_ros2_b200 import orbx, synth
from oracle import pyoracle as o
im = synth.frame(5, 752, 480)
ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
ex(im, None, (0, 1000))
po = o.PortExtractor(nfeatures=1000); po.extract(im, (0, 1000))
for l in range(8):
    c = ex.candidates(0, l); want = po.candidates(l)
    sc = set(map(tuple, c.tolist())); sw = set(map(tuple, want.tolist()))
    extra = sorted(sc - sw, key=lambda t: (t[1], t[0])); missing = sorted(sw - sc, key=lambda t: (t[1], t[0]))
    lw, lh = ex.mvImagePyramid[l].shape[1], ex.mvImagePyramid[l].shape[0]
    width, height = lw - 32, lh - 32
    nC, nR = int(width / 35), int(height / 35)
    wC, hC = int(np.ceil(width / nC)), int(np.ceil(height / nR))
    print(f"level {l}: {lw}x{lh} cells {nR}x{nC} of {hC}x{wC}: mine {len(c)} want {len(want)} extra {len(extra)} missing {len(missing)}")
    for (x, y, r) in extra[:8]:
        j, i = x // wC, y // hC
        print("   extra", (x, y, r), "cell", (i, j), "in-cell", (x - j * wC, y - i * hC), "shift", (35 + j * wC) & 3)
    for (x, y, r) in missing[:8]:
        j, i = x // wC, y // hC
        print("   missing", (x, y, r), "cell", (i, j), "in-cell", (x - j * wC, y - i * hC), "shift", (35 + j * wC) & 3)
