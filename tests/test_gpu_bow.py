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
