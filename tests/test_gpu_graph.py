"""CUDA-graph replay of the fixed-shape pipeline (glue_factory_b200/graph.py) equals the eager pass bit for bit."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _pipe(dev, tmp_path, sp_weights, lg_weights, kpts):
    from gluefactory.models import get_model

    torch.save(sp_weights, tmp_path / "sp.pth")
    torch.save(lg_weights, tmp_path / "lg.pth")
    return get_model("two_view_pipeline")({
        "extractor": {"name": "extractors.superpoint_open", "max_num_keypoints": kpts, "nms_radius": 3,
                      "detection_threshold": -1, "weights": str(tmp_path / "sp.pth")},
        "matcher": {"name": "matchers.lightglue", "filter_threshold": 0.1, "flash": False,
                    "weights": str(tmp_path / "lg.pth")}}).to(dev).eval()


def test_graph_replay_equals_eager(tmp_path, sp_weights, lg_weights):
    from glue_factory_b200 import synth
    from glue_factory_b200.graph import GraphedForward

    dev = torch.device("cuda:0")
    pipe = _pipe(dev, tmp_path, sp_weights, lg_weights, 512)
    pairs = []
    for seed in (3, 4, 5):
        im0, im1, _ = synth.synthetic_pair(seed, 256, 320)  # (1, H, W) each
        pairs.append({"view0": {"image": im0[None].to(dev)}, "view1": {"image": im1[None].to(dev)}})
    with torch.no_grad():
        eager = [{k: v.clone() for k, v in pipe(p).items() if torch.is_tensor(v)} for p in pairs]
    g = GraphedForward(pipe, pairs[0])
    for p, ref in list(zip(pairs, eager)) + [(pairs[0], eager[0])]:
        out = g(p)
        torch.cuda.synchronize()
        for k in ("keypoints0", "keypoints1", "descriptors0", "matches0", "matches1", "matching_scores0",
                  "log_assignment"):
            assert torch.equal(out[k], ref[k]), k
    with pytest.raises(ValueError):
        g({"view0": {"image": pairs[0]["view0"]["image"][..., :300]}, "view1": pairs[0]["view1"]})


def test_graph_refuses_data_dependent_pass(tmp_path, sp_weights, lg_weights):
    """The adaptive matcher reads its final sizes on the host: stream capture must fail loudly, not record garbage."""
    from gluefactory.models import get_model
    from glue_factory_b200 import synth
    from glue_factory_b200.graph import GraphedForward

    dev = torch.device("cuda:0")
    torch.save(lg_weights, tmp_path / "lg.pth")
    m = get_model("matchers.lightglue")({"weights": str(tmp_path / "lg.pth"), "depth_confidence": 0.95,
                                         "width_confidence": 0.95, "filter_threshold": 0.1}).to(dev).eval()
    g = torch.Generator().manual_seed(0)
    data = {"keypoints0": (torch.rand(1, 256, 2, generator=g) * 200).to(dev),
            "keypoints1": (torch.rand(1, 256, 2, generator=g) * 200).to(dev),
            "descriptors0": torch.nn.functional.normalize(torch.randn(1, 256, 256, generator=g), dim=-1).to(dev),
            "descriptors1": torch.nn.functional.normalize(torch.randn(1, 256, 256, generator=g), dim=-1).to(dev)}
    with pytest.raises(Exception):
        GraphedForward(m, data)
    torch.cuda.synchronize()
    with torch.no_grad():
        m(data)  # the library is still usable after the refused capture
