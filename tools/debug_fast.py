import sys, numpy as np
sys.path.insert(0, '.')
from orb_slam3_ros2_b200 import orbx, synth
from oracle import pyoracle as o
im = synth.frame(5, 752, 480)
ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
ex(im, None, (0, 1000))
c = ex.candidates(0, 0)
po = o.PortExtractor(nfeatures=1000); po.extract(im, (0, 1000)); want = po.candidates(0)
wCell, hCell = 36, 38
for (i, j) in [(0, 0), (0, 1), (3, 5)]:
    mine = c[(c[:, 0] // wCell == j) & (c[:, 1] // hCell == i)]
    w = want[(want[:, 0] // wCell == j) & (want[:, 1] // hCell == i)]
    print("cell", i, j, "mine", mine.tolist()[:12], "\n   want", w.tolist()[:12])
    cell = im[16 + i * hCell:16 + i * hCell + hCell + 6, 16 + j * wCell:16 + j * wCell + wCell + 6]
    k = o.fast(cell, 20)
    print("   oracle.fast on cell:", k.tolist()[:12])
    k0 = o.fast(cell, 7, nms=False)
    print("   corners@7 no nms:", len(k0), " @20:", len(o.fast(cell, 20, nms=False)))
