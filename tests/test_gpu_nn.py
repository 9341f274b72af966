"""GPU parity of the nearest-neighbour matcher against the CPU oracle and the
reference-generated golden vectors, through get_model / forward(data) and the C ABI.

Tolerances:
  similarity        |d| <= 5e-5 (two-term bf16 split, fp32 accumulation; unit descriptors)
  matches (teacher-forced on a given fp32 similarity)   bit-exact, ties to the lowest index
  matches (from descriptors)                            agreement >= 0.995 (near-ties of the ratio test may move)
  log_assignment    |d| <= 1e-4 absolute (values ~ -10), teacher-forced
"""
import os

import numpy as np
import pytest
import torch

from oracle import nn_matcher as onn

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch.device("cuda:0")


def load(golden_dir, name):
    with np.load(os.path.join(golden_dir, name + ".npz"), allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def descriptors(seed, b, m, n, d, noise=0.06):
    g = torch.Generator().manual_seed(seed)
    d0 = torch.nn.functional.normalize(torch.randn(b, m, d, generator=g), dim=-1)
    d1 = torch.nn.functional.normalize(torch.randn(b, n, d, generator=g), dim=-1)
    shared = min(m, n) * 2 // 3
    for i in range(b):
        src = torch.randperm(m, generator=g)[:shared]
        dst = torch.randperm(n, generator=g)[:shared]
        d1[i, dst] = torch.nn.functional.normalize(d0[i, src] + noise * torch.randn(shared, d, generator=g), dim=-1)
    return d0, d1


def check_teacher_forced(dev, sim, ratio, dist, mutual):
    """ops.nn_match on a given fp32 similarity against the oracle on the same similarity."""
    from glue_factory_b200 import ops

    want = onn.forward(None, None, ratio, dist, mutual, sim=sim)
    la, m0, m1, ms0, ms1 = ops.nn_match(sim.to(dev), ratio, dist, mutual)
    assert m0.dtype == torch.int64 and m1.dtype == torch.int64
    np.testing.assert_array_equal(m0.cpu().numpy(), want["matches0"].numpy())
    np.testing.assert_array_equal(m1.cpu().numpy(), want["matches1"].numpy())
    np.testing.assert_array_equal(ms0.cpu().numpy(), want["matching_scores0"].numpy())
    np.testing.assert_array_equal(ms1.cpu().numpy(), want["matching_scores1"].numpy())
    np.testing.assert_allclose(la.cpu().numpy(), want["log_assignment"].numpy(), rtol=0, atol=1e-4)
    return want


@pytest.mark.parametrize("name", ["nn_a", "nn_b", "nn_c"])
def test_nn_vs_reference_golden(dev, golden_dir, name):
    from gluefactory.models import get_model

    g = load(golden_dir, name)
    conf = {"ratio_thresh": float(g["ratio_thresh"]) or None, "distance_thresh": float(g["distance_thresh"]) or None,
            "mutual_check": bool(g["mutual_check"])}
    # stage 1: decisions on the reference's own similarity are bit-exact
    sim_ref = torch.from_numpy(g["out_similarity"])
    from glue_factory_b200 import ops

    la, m0, m1, ms0, ms1 = ops.nn_match(sim_ref.to(dev), conf["ratio_thresh"], conf["distance_thresh"], conf["mutual_check"])
    np.testing.assert_array_equal(m0.cpu().numpy(), g["out_matches0"])
    np.testing.assert_array_equal(m1.cpu().numpy(), g["out_matches1"])
    np.testing.assert_array_equal(ms0.cpu().numpy(), g["out_matching_scores0"])
    np.testing.assert_array_equal(ms1.cpu().numpy(), g["out_matching_scores1"])
    np.testing.assert_allclose(la.cpu().numpy(), g["out_log_assignment"], rtol=0, atol=1e-4)
    # stage 2: the model from descriptors
    model = get_model("matchers.nearest_neighbor_matcher")(conf).to(dev).eval()
    with torch.no_grad():
        pred = model({"descriptors0": torch.from_numpy(g["descriptors0"]).to(dev),
                      "descriptors1": torch.from_numpy(g["descriptors1"]).to(dev)})
    assert set(pred) == {"matches0", "matches1", "matching_scores0", "matching_scores1", "similarity", "log_assignment"}
    err = np.abs(pred["similarity"].cpu().numpy() - g["out_similarity"]).max()
    agree0 = (pred["matches0"].cpu().numpy() == g["out_matches0"]).mean()
    agree1 = (pred["matches1"].cpu().numpy() == g["out_matches1"]).mean()
    print(f"{name}: similarity max |d| = {err:.2e}; match agreement {agree0:.4f} / {agree1:.4f}")
    assert err < 5e-5
    assert agree0 >= 0.995 and agree1 >= 0.995
    np.testing.assert_allclose(pred["log_assignment"].cpu().numpy(), g["out_log_assignment"], rtol=0, atol=3e-4)


@pytest.mark.parametrize("B,M,N,ratio,dist,mutual", [
    (1, 1, 1, None, None, True),
    (2, 1, 37, None, 0.8, True),
    (1, 64, 256, 0.9, None, True),
    (3, 65, 257, 0.95, 1.2, False),
    (2, 300, 131, 0.8, None, True),
    (1, 513, 1030, None, None, False),
    (2, 2048, 2048, 0.9, 1.1, True),
])
def test_nn_match_teacher_forced(dev, B, M, N, ratio, dist, mutual):
    d0, d1 = descriptors(100 + M + N, B, M, N, 64)
    if ratio and min(M, N) < 2:
        pytest.skip("ratio test needs two candidates")
    check_teacher_forced(dev, onn.similarity(d0, d1), ratio, dist, mutual)


def test_nn_match_ties_take_lowest_index(dev):
    sim = torch.zeros(1, 70, 300)
    sim[0, :, 5] = 0.5
    sim[0, :, 260] = 0.5   # equal maximum in another 256-column tile
    sim[0, 3, :] = 0.25
    sim[0, 66, :] = 0.25   # equal column maximum in the second 32-row half
    sim[0, 3, 5] = sim[0, 66, 5] = 0.75
    want = check_teacher_forced(dev, sim, None, None, False)
    assert want["matches0"][0, 0].item() == 5 and want["matches1"][0, 5].item() == 3


def test_nn_similarity_accuracy(dev):
    from glue_factory_b200.models.nearest_neighbor_matcher import NearestNeighborMatcher

    d0, d1 = descriptors(7, 2, 700, 900, 256)
    sim = NearestNeighborMatcher.similarity(d0.to(dev), d1.to(dev)).cpu()
    want = (d0.double() @ d1.double().transpose(1, 2)).float()
    err = (sim - want).abs().max().item()
    print("similarity (D = 256) max |d| =", err)
    assert err < 3e-5


def test_nn_full_size_properties(dev):
    """BASELINE size (32 pairs x 2048 x 2048): size-independent properties instead of the oracle."""
    from gluefactory.models import get_model

    B, M, N = 32, 2048, 2048
    d0, d1 = descriptors(5, B, M, N, 256)
    model = get_model("matchers.nearest_neighbor_matcher")({"ratio_thresh": 0.9}).to(dev).eval()
    with torch.no_grad():
        pred = model({"descriptors0": d0.to(dev), "descriptors1": d1.to(dev)})
    m0, m1 = pred["matches0"], pred["matches1"]
    assert m0.shape == (B, M) and m1.shape == (B, N)
    valid = m0 > -1
    assert 0.3 < valid.float().mean().item() < 0.9
    back = torch.gather(m1, 1, m0.clamp(min=0))
    assert torch.equal(back[valid], torch.arange(M, device=dev).expand(B, M)[valid])   # mutual consistency
    assert int(valid.sum()) == int((m1 > -1).sum())
    assert torch.equal(pred["matching_scores0"], valid.float())
    # a matched pair is the row arg-max of the similarity
    sim = pred["similarity"]
    assert torch.equal(sim.argmax(-1)[valid], m0[valid])
    la = pred["log_assignment"]
    assert la.shape == (B, M + 1, N + 1)
    assert la[:, -1].abs().max().item() == 0 and la[:, :, -1].abs().max().item() == 0
    # exp of each log-softmax sums to one: logsumexp over a row of (la - column term) -- checked on one pair
    row_ls = torch.log_softmax(sim[0], -1) + torch.log_softmax(sim[0], -2)
    assert (la[0, :M, :N] - row_ls).abs().max().item() < 1e-4


def test_nn_rejects_unsupported(dev):
    from glue_factory_b200._lib import B200Error
    from gluefactory.models import get_model

    model = get_model("matchers.nearest_neighbor_matcher")({}).eval()
    with pytest.raises(B200Error):
        model({"descriptors0": torch.randn(1, 4, 64), "descriptors1": torch.randn(1, 4, 64)})  # CPU tensors
    with pytest.raises(NotImplementedError):
        model({"descriptors0": torch.randn(1, 4, 48, device=dev), "descriptors1": torch.randn(1, 4, 48, device=dev)})
    with pytest.raises(AssertionError):
        model({"descriptors0": torch.randn(1, 4, 64, device=dev)})  # missing key, as BaseModel.forward reports it
    with pytest.raises(NotImplementedError):
        model.loss({}, {})


def test_superpoint_open_nn_pipeline(dev, tmp_path, sp_weights):
    """configs/eval/superpoint-open+NN.yaml: SP-open extractor + mutual NN matcher through TwoViewPipeline."""
    from glue_factory_b200 import synth
    from gluefactory.models import get_model

    torch.save(sp_weights, tmp_path / "sp.pth")
    conf = {"extractor": {"name": "extractors.superpoint_open", "weights": str(tmp_path / "sp.pth"),
                          "max_num_keypoints": 512, "detection_threshold": 0.0, "nms_radius": 3},
            "matcher": {"name": "matchers.nearest_neighbor_matcher"}}
    pipe = get_model("two_view_pipeline")(conf).to(dev).eval()
    img0, img1, _ = synth.synthetic_pair(3, 256, 320)
    data = {"view0": {"image": img0[None].to(dev)}, "view1": {"image": img1[None].to(dev)}}
    with torch.no_grad():
        pred = pipe(data)
    want = onn.forward(pred["descriptors0"].cpu(), pred["descriptors1"].cpu())
    agree = (pred["matches0"].cpu() == want["matches0"]).float().mean().item()
    n_match = int((pred["matches0"] > -1).sum())
    print(f"SP-open + NN pipeline: {n_match} mutual matches, agreement with the oracle {agree:.4f}")
    assert agree >= 0.995 and n_match > 50
