"""GPU parity (through the C ABI) of row N4, Frame::ComputeBoW: bowDescendKernel + bowCollectKernel against the
reference DBoW2 fixture (tests/golden/bow.npz) and the CPU oracle; doubles bit for bit."""
import numpy as np
import pytest

from conftest import load_golden
from orb_slam3_ros2_b200 import synth
from test_oracle_bow import CASES, assert_same, golden_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orbx():
    from orb_slam3_ros2_b200 import orbx as m
    m.lib()
    return m


@pytest.mark.parametrize("name", CASES)
def test_transform_matches_reference_dbow2_fixture(orbx, oracle, name):
    g = load_golden("bow.npz")
    k, L, levelsup, parent, node_desc, weight, desc = synth.bow_case(name)
    voc = orbx.ORBVocabulary(parent, node_desc, weight, L)
    info = voc.info()
    assert info["n_nodes"] == len(parent) and info["n_words"] == int(g[f"{name}__words"]) and info["L"] == L
    assert_same(voc.transform(desc, levelsup), golden_case(g, name), name)
    for lu in (0, L):
        if name == "k10L4_earlyleaf" and L - lu > 3:
            continue
        assert_same(voc.transform(desc, lu), oracle.bow_transform(parent, node_desc, weight, L, desc, lu), (name, lu))
    voc.close()


def test_batch_of_ragged_frames_and_empty_frame(orbx, oracle):
    k, L, levelsup, parent, node_desc, weight, desc = synth.bow_case("k10L4_levelsup2")
    voc = orbx.ORBVocabulary(parent, node_desc, weight, L)
    frames = [desc[:700], desc[700:701], desc[:0], desc[300:2000], synth.descriptors_near_words(parent, node_desc, 1234, seed=3)]
    got = voc.transform_batch(frames, levelsup)
    for f, d in enumerate(frames):
        assert_same(got[f], oracle.bow_transform(parent, node_desc, weight, L, d, levelsup), f)
    assert len(got[2]["word_ids"]) == 0 and len(got[2]["node_ids"]) == 0 and list(got[2]["node_begin"]) == [0]


def test_all_words_stopped_and_duplicate_descriptors(orbx, oracle):
    k, L, levelsup, parent, node_desc, weight, desc = synth.bow_case("k10L3")
    # every feature identical: one word, one node, value 1.0, all indices listed
    same = np.repeat(desc[:1], 500, axis=0)
    voc = orbx.ORBVocabulary(parent, node_desc, weight, L)
    out = voc.transform(same, levelsup)
    assert_same(out, oracle.bow_transform(parent, node_desc, weight, L, same, levelsup), "same")
    assert len(out["word_ids"]) == 1 and out["word_vals"][0] == 1.0 and list(out["feat_idx"]) == list(range(500))
    # all weights zero: both maps stay empty (w > 0 is the "not stopped" test)
    stopped = orbx.ORBVocabulary(parent, node_desc, np.zeros_like(weight), L)
    out = stopped.transform(desc, levelsup)
    assert_same(out, oracle.bow_transform(parent, node_desc, np.zeros_like(weight), L, desc, levelsup), "stopped")
    assert len(out["word_ids"]) == 0 and len(out["feat_idx"]) == 0


def test_transform_on_device_resident_descriptors_of_an_extraction(orbx, oracle):
    g = load_golden("bow.npz")
    k, L, levelsup, parent, node_desc, weight, _ = synth.bow_case("k10L4_levelsup2")
    voc = orbx.ORBVocabulary(parent, node_desc, weight, L)
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    # the first frame of the batch is the e2e_euroc752 fixture
    gold = load_golden("e2e_euroc752.npz")
    imgs = synth.frames(int(gold["k"]), 3, 752, 480)
    ex.extract_batch(imgs, (0, 1000), download=False)
    got = voc.transform_extracted(ex, levelsup)
    assert len(got) == 3
    assert_same(got[0], golden_case(g, "euroc752"), "euroc752 (device-resident)")
    for f in range(3):
        _, _, desc = ex.download(f)
        assert_same(got[f], oracle.bow_transform(parent, node_desc, weight, L, desc, levelsup), f)


def test_bad_arguments(orbx):
    with pytest.raises(orbx.OrbxError):
        orbx.ORBVocabulary(np.array([-1, 0, 3, 1], np.int32), np.zeros((4, 32), np.uint8), np.ones(4), 2)


@pytest.mark.parametrize("seed,levelsup,nnratio,ori", [(0, 2, 0.7, True), (1, 2, 0.9, True), (2, 3, 0.7, False), (3, 1, 0.6, True),
                                                      (4, 4, 0.75, True)])
def test_search_by_bow_matches_oracle(orbx, oracle, seed, levelsup, nnratio, ori):
    """ORBmatcher::SearchByBoW(KF, F): the oracle is pinned to the reference's own body in tests/test_oracle_refslices.py."""
    from test_oracle_refslices import bow_pair_case
    case = bow_pair_case(oracle, seed, 1000, levelsup)
    n_want, m_want = oracle.search_by_bow(*case, nnratio=nnratio, check_orientation=ori)
    n_got, m_got = orbx.search_by_bow(*case, nnratio=nnratio, check_orientation=ori)
    assert n_got == n_want and np.array_equal(m_got, m_want)
    if oracle.have_refslices():
        n_ref, m_ref = oracle.ref_search_by_bow(*case, nnratio=nnratio, check_orientation=ori)
        assert n_got == n_ref and np.array_equal(m_got, m_ref)


def test_search_by_bow_chain_on_extracted_frames(orbx, oracle):
    """extract -> ComputeBoW (device-resident descriptors) -> SearchByBoW between consecutive frames, all on the device path."""
    k, L, levelsup, parent, node_desc, weight, _ = synth.bow_case("k10L4_levelsup2")
    voc = orbx.ORBVocabulary(parent, node_desc, weight, L)
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    ex.extract_batch(synth.frames(0, 2, 752, 480), (0, 1000), download=False)
    fv = voc.transform_extracted(ex, levelsup)
    (_, k0, d0), (_, k1, d1) = ex.download(0), ex.download(1)
    valid = np.ones(len(k0), np.uint8)
    got = orbx.search_by_bow(d0, valid, k0["angle"], fv[0], d1, k1["angle"], fv[1], 0.7, True)
    want = oracle.search_by_bow(d0, valid, k0["angle"], fv[0], d1, k1["angle"], fv[1], 0.7, True)
    assert got[0] == want[0] and np.array_equal(got[1], want[1])
    # empty keyframe side / nothing valid
    n0, m0 = orbx.search_by_bow(d0, np.zeros(len(k0), np.uint8), k0["angle"], fv[0], d1, k1["angle"], fv[1])
    assert n0 == 0 and (m0 == -1).all()


@pytest.mark.parametrize("seed,levelsup,nnratio,ori", [(0, 2, 0.75, True), (1, 2, 0.9, True), (2, 3, 0.75, False), (3, 1, 0.6, True),
                                                      (4, 4, 0.75, True)])
def test_search_by_bow_keyframes_matches_oracle(orbx, oracle, seed, levelsup, nnratio, ori):
    """ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, ...) (ORBmatcher.cc:765-905) on the device == oracle == the reference's body."""
    from test_oracle_refslices import bow_kf_pair_case
    case = bow_kf_pair_case(oracle, seed, 1000, levelsup)
    n_want, m_want = oracle.search_by_bow_kf(*case, nnratio=nnratio, check_orientation=ori)
    n_got, m_got = orbx.search_by_bow_kf(*case, nnratio=nnratio, check_orientation=ori)
    assert n_got == n_want and np.array_equal(m_got, m_want)
    if oracle.have_refslices():
        n_ref, m_ref = oracle.ref_search_by_bow_kf(*case, nnratio=nnratio, check_orientation=ori)
        assert n_got == n_ref and np.array_equal(m_got, m_ref)
    # nothing valid on the second side
    d1, v1, a1, fv1, d2, v2, a2, fv2 = case
    n0, m0 = orbx.search_by_bow_kf(d1, v1, a1, fv1, d2, np.zeros_like(v2), a2, fv2, nnratio, ori)
    assert n0 == 0 and (m0 == -1).all()
    # distance exactly TH_LOW is rejected here (strict, :848) but accepted by the keyframe -> frame form (:323)
    q = np.zeros((1, 32), np.uint8)
    t = np.zeros((1, 32), np.uint8); t[0, :6] = 0xff; t[0, 6] = 0x03  # 50 bits
    fv = {"node_ids": np.array([7], np.int32), "node_begin": np.array([0, 1], np.int32), "feat_idx": np.array([0], np.int32)}
    one, ang = np.ones(1, np.uint8), np.zeros(1, np.float32)
    assert orbx.search_by_bow_kf(q, one, ang, fv, t, one, ang, fv, 0.75, False)[0] == 0
    assert orbx.search_by_bow(q, one, ang, fv, t, ang, fv, 0.75, False)[0] == 1
