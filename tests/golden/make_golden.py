"""Generate golden vectors by running the UNMODIFIED reference (authoring container only).

    python tests/golden/make_golden.py

Imports `gluefactory` from /root/reference (with the omegaconf stand-in under oracle/_shim,
because omegaconf is not installable here) and stores small input/output fixtures as .npz
under tests/golden/.  The fixtures pin the CPU oracle (tests/test_oracle_golden.py) and,
through it, the CUDA path.  /root/reference does not exist on the GPU box; nothing at test
time imports it.

Deviation recorded for the adaptive case: the reference's adaptive branch raises as shipped
(lightglue.py:518,538,550), so the instance gets `confidence_thresholds` from its own
`confidence_threshold()` formula and the empty `ref_descriptors` stack is guarded.
"""
import os
import sys
import tempfile
import warnings

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != REPO]
sys.path[:0] = ["/root/reference", os.path.join(REPO, "oracle", "_shim")]
sys.path.append(REPO)  # only for glue_factory_b200.synth / _conf (no `gluefactory` shadowing: reference is first)

import cv2  # noqa: E402
import numpy as np  # noqa: E402
import torch  # noqa: E402

warnings.filterwarnings("ignore")
import gluefactory  # noqa: E402

assert gluefactory.__file__.startswith("/root/reference"), gluefactory.__file__
from gluefactory.models import get_model  # noqa: E402

from glue_factory_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
torch.set_num_threads(8)
torch.manual_seed(0)


def save(name, **arrays):
    arrays = {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in arrays.items()}
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **arrays)
    print(name, {k: v.shape for k, v in arrays.items()})


def quantised_image(seed, h, w):
    img = synth.synthetic_image(seed, h, w)
    return torch.round(img * 255).to(torch.uint8)


def main():
    sp_sd = synth.superpoint_open_state_dict()
    lg_sd = synth.lightglue_state_dict()
    tmp = tempfile.mkdtemp()
    sp_path = os.path.join(tmp, "sp.pth")
    lg_path = os.path.join(tmp, "lg.pth")
    torch.save(sp_sd, sp_path)
    torch.save(lg_sd, lg_path)
    with open(os.path.join(OUT, "weights.sha256"), "w") as fid:
        fid.write(f"superpoint_open {synth.state_dict_sha256(sp_sd)}\nlightglue {synth.state_dict_sha256(lg_sd)}\n")

    SP = get_model("extractors.superpoint_open")
    LG = get_model("matchers.lightglue")

    # ---- 1. SuperPoint-open, two configurations, one image each -------------------------
    for tag, (h, w), conf in (
        ("sp_a", (120, 160), dict(nms_radius=4, max_num_keypoints=96, detection_threshold=0.005)),
        ("sp_b", (136, 104), dict(nms_radius=3, max_num_keypoints=64, detection_threshold=-1)),
        ("sp_c", (96, 128), dict(nms_radius=2, max_num_keypoints=None, detection_threshold=0.02, remove_borders=6)),
    ):
        model = SP({**conf, "weights": sp_path}).eval()
        img_u8 = quantised_image(77 + h, h, w)
        with torch.no_grad():
            pred = model({"image": img_u8[None].float() / 255.0})
        save(
            tag,
            image_u8=img_u8,
            conf=np.array(repr(conf)),
            keypoints=pred["keypoints"],
            keypoint_scores=pred["keypoint_scores"],
            descriptors=pred["descriptors"],
            heatmap=pred["heatmap"],
        )

    # RGB input path
    model = SP({"max_num_keypoints": 32, "weights": sp_path}).eval()
    rgb = torch.stack([quantised_image(300 + c, 64, 80)[0] for c in range(3)])
    with torch.no_grad():
        pred = model({"image": rgb[None].float() / 255.0})
    save("sp_rgb", image_u8=rgb, keypoints=pred["keypoints"], keypoint_scores=pred["keypoint_scores"],
         descriptors=pred["descriptors"])

    # ---- 2. LightGlue, non-adaptive, batch 2, M != N -------------------------------------
    g = np.random.default_rng(5)
    m, n = 96, 80
    size = np.array([[320.0, 240.0], [320.0, 240.0]], dtype=np.float32)
    k0 = (g.random((2, m, 2)) * size[:, None, :]).astype(np.float32)
    k1 = (g.random((2, n, 2)) * size[:, None, :]).astype(np.float32)
    d0 = g.standard_normal((2, m, 256)).astype(np.float32)
    d1 = np.concatenate([d0[:, :n] + 0.35 * g.standard_normal((2, n, 256)).astype(np.float32)], 1)
    d0 /= np.linalg.norm(d0, axis=-1, keepdims=True)
    d1 /= np.linalg.norm(d1, axis=-1, keepdims=True)
    lg = LG({"weights": lg_path, "filter_threshold": 0.1}).eval()
    data = {
        "keypoints0": torch.from_numpy(k0), "keypoints1": torch.from_numpy(k1),
        "descriptors0": torch.from_numpy(d0), "descriptors1": torch.from_numpy(d1),
        "view0": {"image_size": torch.from_numpy(size)}, "view1": {"image_size": torch.from_numpy(size)},
    }
    with torch.no_grad():
        pred = lg(data)
    save("lg_a", keypoints0=k0, keypoints1=k1, descriptors0=d0, descriptors1=d1, image_size=size,
         **{k: v for k, v in pred.items()})
    # same, without image_size (extent-based normalisation)
    data2 = {**data, "view0": {}, "view1": {}}
    with torch.no_grad():
        pred = lg(data2)
    save("lg_nosize", log_assignment=pred["log_assignment"], matches0=pred["matches0"], matches1=pred["matches1"])

    # ---- 3. LightGlue adaptive (patched as documented above), batch 1 -------------------
    m = n = 160
    k0 = (g.random((1, m, 2)) * size[:1, None, :]).astype(np.float32)
    k1 = (g.random((1, n, 2)) * size[:1, None, :]).astype(np.float32)
    d0 = g.standard_normal((1, m, 256)).astype(np.float32)
    d1 = d0[:, g.permutation(m)] + 0.35 * g.standard_normal((1, n, 256)).astype(np.float32)
    d0 /= np.linalg.norm(d0, axis=-1, keepdims=True)
    d1 /= np.linalg.norm(d1, axis=-1, keepdims=True)
    for tag, dc, wc in (("lg_adaptive", 0.95, 0.95), ("lg_prune_only", -1, 0.9), ("lg_depth_only", 0.55, -1)):
        lg = LG({"weights": lg_path, "filter_threshold": 0.1, "depth_confidence": dc, "width_confidence": wc}).eval()
        lg.confidence_thresholds = [lg.confidence_threshold(i) for i in range(lg.conf.n_layers)]
        data = {
            "keypoints0": torch.from_numpy(k0), "keypoints1": torch.from_numpy(k1),
            "descriptors0": torch.from_numpy(d0), "descriptors1": torch.from_numpy(d1),
            "view0": {"image_size": torch.from_numpy(size[:1])}, "view1": {"image_size": torch.from_numpy(size[:1])},
        }
        orig_stack = torch.stack
        last = {}
        orig_layers = [t.forward for t in lg.transformers]

        def make_hook(fwd):
            def hooked(*a, **kw):
                out = fwd(*a, **kw)
                last["d"] = out
                return out
            return hooked

        for t in lg.transformers:
            t.forward = make_hook(t.forward)

        def guarded_stack(tensors, *a, **kw):
            if len(tensors) == 0:
                idx = guarded_stack.calls
                guarded_stack.calls += 1
                return orig_stack([last["d"][idx]], *a, **kw)
            return orig_stack(tensors, *a, **kw)

        guarded_stack.calls = 0
        torch.stack = guarded_stack
        try:
            with torch.no_grad():
                pred = lg(data)
        finally:
            torch.stack = orig_stack
        save(tag, keypoints0=k0, keypoints1=k1, descriptors0=d0, descriptors1=d1, image_size=size[:1],
             depth_confidence=dc, width_confidence=wc, **{k: v for k, v in pred.items()})

    # ---- 4. configuration 1: boat pair through TwoViewPipeline, 1024 keypoints ----------
    pipe = get_model("two_view_pipeline")(
        {
            "extractor": {"name": "extractors.superpoint_open", "max_num_keypoints": 1024, "nms_radius": 4,
                          "detection_threshold": 0.0, "weights": sp_path},
            "matcher": {"name": "matchers.lightglue", "filter_threshold": 0.1, "weights": lg_path},
        }
    ).eval()
    views = {}
    for i, name in enumerate(("boat1.png", "boat2.png")):
        im = cv2.imread(os.path.join(OUT, "assets", name), cv2.IMREAD_GRAYSCALE)
        t = torch.from_numpy(im)[None, None].float() / 255.0
        views[f"view{i}"] = {"image": t, "image_size": torch.tensor([[im.shape[1], im.shape[0]]]).float()}
    with torch.no_grad():
        pred = pipe(views)
    la = pred["log_assignment"]
    save(
        "boat_pipeline",
        keypoints0=pred["keypoints0"], keypoints1=pred["keypoints1"],
        keypoint_scores0=pred["keypoint_scores0"], keypoint_scores1=pred["keypoint_scores1"],
        descriptors0_head=pred["descriptors0"][:, :48], descriptors1_head=pred["descriptors1"][:, :48],
        matches0=pred["matches0"], matches1=pred["matches1"],
        matching_scores0=pred["matching_scores0"], matching_scores1=pred["matching_scores1"],
        log_assignment_sub=la[:, ::8, ::8], log_assignment_rowmax=la[:, :-1, :-1].max(2).values,
        heatmap0_sub=pred["heatmap0"][:, ::4, ::4],
    )


if __name__ == "__main__":
    main()
