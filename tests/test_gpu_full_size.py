"""BASELINE.json's configs at FULL size on the GPU, through size-independent properties (the oracle only sees samples):
  config 2   256 frames 640x480 / 1000 features + frame-to-frame top-2: batch == any other cut of the same stream,
             sampled frames == oracle, every match distance recomputed from the descriptors;
  config 4   10,000 frames 1920x1080 / 4000 features generated on the device: two passes with different chunkings
             (250 frames on one stream pair vs 96 frames in 2 sub-batches) give identical sampled frames and identical
             counts for ALL frames; sampled frames == oracle.
Frames are independent in the reference (ORBextractor keeps no state between calls), which is what these cuts test."""
import numpy as np
import pytest

from orb_slam3_ros2_b200 import synth

pytestmark = pytest.mark.gpu

LAP = (0, 1000)


@pytest.fixture(scope="module")
def orbx():
    from orb_slam3_ros2_b200 import orbx as m
    m.lib()
    return m


def run_stream(orbx, ex, total, chunk, w, h, sample_every):
    """Extract frames 0..total-1 of the synthetic stream in chunks; -> (counts of all frames, {k: (mono, kps, desc)})."""
    import torch
    buf = torch.empty((chunk, h, w), dtype=torch.uint8, device="cuda")
    counts = np.zeros(total, np.int32)
    sampled = {}
    for k0 in range(0, total, chunk):
        F = min(chunk, total - k0)
        orbx._check(orbx.lib().orbx_synth_frames_device(buf.data_ptr(), k0, F, w, h, 42, 0, -1))
        torch.cuda.synchronize()
        ex.extract_batch_device(buf.data_ptr(), F, w, h, LAP)
        ex.sync()
        counts[k0:k0 + F] = ex.counts()[0]  # all frames of the chunk in one copy
        for f in range(F):
            if (k0 + f) % sample_every == 0:
                sampled[k0 + f] = ex.download(f)
    return counts, sampled


def same_frame(a, b):
    return a[0] == b[0] and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])


def test_config4_10k_frames_1080p_chunking_invariance(orbx, oracle):
    total, w, h, nf = 10_000, 1920, 1080, 4000
    ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    counts_a, samp_a = run_stream(orbx, ex, total, 250, w, h, 125)
    ex.set_subbatches(2)
    counts_b, samp_b = run_stream(orbx, ex, total, 96, w, h, 125)
    assert np.array_equal(counts_a, counts_b)
    assert counts_a.min() > nf * 0.9 and counts_a.max() <= ex.capacity
    assert sorted(samp_a) == sorted(samp_b) and len(samp_a) == total // 125
    for k in samp_a:
        assert same_frame(samp_a[k], samp_b[k]), k
        assert len(samp_a[k][1]) == counts_a[k]
    po = oracle.PortExtractor(nfeatures=nf)
    for k in (0, 9875):
        pm, pk, pd = po.extract(synth.frame(k, w, h), LAP)
        assert same_frame(samp_a[k], (pm, pk, pd)), k


def test_config2_256_frames_batch_and_matches(orbx, oracle):
    import torch
    F, w, h, nf = 256, 640, 480, 1000
    ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    ex.set_subbatches(2)
    counts, samp = run_stream(orbx, ex, F, F, w, h, 1)          # one batch of 256
    i1, d1, i2, d2 = ex.match_consecutive()                     # 255 frame pairs
    ex2 = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    counts_c, samp_c = run_stream(orbx, ex2, F, 37, w, h, 1)    # the same stream cut into 37-frame batches
    assert np.array_equal(counts, counts_c)
    for k in range(F):
        assert same_frame(samp[k], samp_c[k]), k
    po = oracle.PortExtractor(nfeatures=nf)
    for k in (0, 131, 255):
        assert same_frame(samp[k], po.extract(synth.frame(k, w, h), LAP)), k
    # every reported distance is the Hamming distance of the pair it names; best <= second; oracle on sampled pairs
    pop = np.unpackbits(np.arange(256, dtype=np.uint8)[:, None], axis=1).sum(1).astype(np.int32)
    for f in range(F - 1):
        n = counts[f]
        a, b = samp[f][2], samp[f + 1][2]
        j1, j2 = i1[f, :n], i2[f, :n]
        assert j1.min() >= 0 and j1.max() < counts[f + 1] and j2.max() < counts[f + 1]
        assert np.array_equal(pop[a ^ b[j1]].sum(1), d1[f, :n])
        ok2 = j2 >= 0
        assert np.array_equal(pop[a[ok2] ^ b[j2[ok2]]].sum(1), d2[f, :n][ok2])
        assert np.all(d1[f, :n] <= d2[f, :n])
        if f in (0, 100, 254):
            want = oracle.match_top2(a, b)
            for g_, w_ in zip((j1, d1[f, :n], j2, d2[f, :n]), want):
                assert np.array_equal(g_, w_)
