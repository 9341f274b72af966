"""Launch state of the library is per device and the models make their tensors' device current: a
model moved to cuda:1 works while cuda:0 is the current device, and equals the cuda:0 result bit for bit.
Skipped on single-GPU boxes."""
import pytest
import torch

from glue_factory_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_pipeline_on_the_non_current_device(tmp_path, sp_weights, lg_weights):
    from glue_factory_b200 import ops
    from glue_factory_b200._lib import B200Error
    from gluefactory.models import get_model

    torch.save(sp_weights, tmp_path / "sp.pth")
    torch.save(lg_weights, tmp_path / "lg.pth")
    conf = {"extractor": {"name": "extractors.superpoint_open", "weights": str(tmp_path / "sp.pth"), "max_num_keypoints": 256,
                          "detection_threshold": -1, "nms_radius": 3},
            "matcher": {"name": "matchers.lightglue", "weights": str(tmp_path / "lg.pth"), "filter_threshold": 0.1}}
    im0, im1, _ = synth.synthetic_batch(3, 2, 160, 192)
    size = torch.tensor([[192.0, 160.0]] * 2)
    out = []
    torch.cuda.set_device(0)
    for d in (0, 1):
        dev = torch.device("cuda", d)
        pipe = get_model("two_view_pipeline")(conf).to(dev).eval()
        with torch.no_grad():  # the current device stays cuda:0 for both
            pred = pipe({"view0": {"image": im0.to(dev), "image_size": size.to(dev)},
                         "view1": {"image": im1.to(dev), "image_size": size.to(dev)}})
        assert pred["matches0"].device == dev and torch.cuda.current_device() == 0
        out.append({k: v.cpu() for k, v in pred.items() if torch.is_tensor(v)})
    for k in ("keypoints0", "keypoint_scores1", "descriptors0", "matches0", "matching_scores1", "log_assignment"):
        assert torch.equal(out[0][k], out[1][k]), k
    # raw wrappers refuse tensors of a device that is not current instead of launching on the wrong one
    with pytest.raises(B200Error):
        ops.sp_nms_border(torch.zeros(1, 32, 32, device="cuda:1"), 2, 0)
