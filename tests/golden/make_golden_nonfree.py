"""Generate golden vectors for the SuperGlue configuration (BASELINE.json configs[3]) by running the
UNMODIFIED reference modules `gluefactory_nonfree.superpoint` / `gluefactory_nonfree.superglue`
(authoring container only).

    python tests/golden/make_golden_nonfree.py

Same mechanics as make_golden.py: the reference is imported from /root/reference with the omegaconf
stand-in; `torch.hub.load_state_dict_from_url` is patched to hand the seeded state dict to SuperPoint
(which otherwise always downloads, superpoint.py:198-200); SuperGlue is built with `weights: None`
and loaded with the seeded state dict.  Nothing at test time imports /root/reference.
"""
import os
import sys
import warnings

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != REPO]
sys.path[:0] = ["/root/reference", os.path.join(REPO, "oracle", "_shim")]
sys.path.append(REPO)

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
    return torch.round(synth.synthetic_image(seed, h, w) * 255).to(torch.uint8)


def main():
    sp_sd = synth.superpoint_nonfree_state_dict()
    sg_sd = synth.superglue_state_dict()
    with open(os.path.join(OUT, "weights_nonfree.sha256"), "w") as fid:
        fid.write(f"superpoint_nonfree {synth.state_dict_sha256(sp_sd)}\nsuperglue {synth.state_dict_sha256(sg_sd)}\n")
    torch.hub.load_state_dict_from_url = lambda *a, **k: sp_sd
    SP = get_model("gluefactory_nonfree.superpoint")
    SG = get_model("gluefactory_nonfree.superglue")

    # ---- SuperPoint (MagicLeap): legacy / fixed sampling, top-k / all candidates, per-image border ----
    for tag, (h, w), conf, size in (
        ("spn_a", (120, 160), dict(nms_radius=4, max_num_keypoints=96, detection_threshold=0.005), None),
        ("spn_b", (136, 104), dict(nms_radius=3, max_num_keypoints=64, detection_threshold=0.0, legacy_sampling=False), None),
        ("spn_c", (96, 128), dict(nms_radius=2, max_num_keypoints=-1, detection_threshold=0.08, remove_borders=6), None),
        ("spn_d", (128, 160), dict(nms_radius=3, max_num_keypoints=80, detection_threshold=0.005), (150.0, 117.0)),
    ):
        model = SP(conf).eval()
        img_u8 = quantised_image(177 + h, h, w)
        data = {"image": img_u8[None].float() / 255.0}
        if size is not None:
            data["image_size"] = torch.tensor([size])
        with torch.no_grad():
            pred = model(data)
        save(tag, image_u8=img_u8, conf=np.array(repr(conf)), image_size=np.array(size if size else (0.0, 0.0), dtype=np.float32),
             keypoints=pred["keypoints"], keypoint_scores=pred["keypoint_scores"], descriptors=pred["descriptors"])

    # ---- SuperGlue: batch 2, M != N, 20 Sinkhorn iterations; and the default 18 layers / 50 iterations ----
    g = np.random.default_rng(11)
    m, n = 96, 80
    size = np.array([[320.0, 240.0], [320.0, 240.0]], dtype=np.float32)
    k0 = (g.random((2, m, 2)) * size[:, None, :]).astype(np.float32)
    k1 = (g.random((2, n, 2)) * size[:, None, :]).astype(np.float32)
    d0 = g.standard_normal((2, m, 256)).astype(np.float32)
    d1 = d0[:, :n] + 0.35 * g.standard_normal((2, n, 256)).astype(np.float32)
    d0 /= np.linalg.norm(d0, axis=-1, keepdims=True)
    d1 /= np.linalg.norm(d1, axis=-1, keepdims=True)
    s0, s1 = g.random((2, m)).astype(np.float32), g.random((2, n)).astype(np.float32)
    img = torch.zeros(2, 1, 240, 320)
    for tag, conf, with_size in (("sg_a", dict(num_sinkhorn_iterations=20, filter_threshold=0.2), True),
                                 ("sg_b", dict(num_sinkhorn_iterations=50, filter_threshold=0.05), False)):
        sg = SG({"weights": None, **conf}).eval()
        sg.load_state_dict(sg_sd)
        view = {"image": img, "image_size": torch.from_numpy(size)} if with_size else {"image": img}
        data = {"keypoints0": torch.from_numpy(k0), "keypoints1": torch.from_numpy(k1),
                "descriptors0": torch.from_numpy(d0), "descriptors1": torch.from_numpy(d1),
                "keypoint_scores0": torch.from_numpy(s0), "keypoint_scores1": torch.from_numpy(s1),
                "view0": view, "view1": view}
        with torch.no_grad():
            pred = sg(data)
        save(tag, keypoints0=k0, keypoints1=k1, descriptors0=d0, descriptors1=d1, scores0=s0, scores1=s1,
             image_size=size, conf=np.array(repr(conf)), **{k: v for k, v in pred.items()})

    # ---- configuration 4 in small: SuperPoint -> SuperGlue through TwoViewPipeline --------------------
    pipe = get_model("two_view_pipeline")({
        "extractor": {"name": "gluefactory_nonfree.superpoint", "max_num_keypoints": 256, "nms_radius": 3,
                      "detection_threshold": 0.0},
        "matcher": {"name": "gluefactory_nonfree.superglue", "weights": None, "num_sinkhorn_iterations": 100,
                    "filter_threshold": 0.2},
    }).eval()
    pipe.matcher.load_state_dict(sg_sd)
    im0, im1, _ = synth.synthetic_pair(5, 192, 256)
    u0, u1 = torch.round(im0 * 255).to(torch.uint8), torch.round(im1 * 255).to(torch.uint8)
    size = torch.tensor([[256.0, 192.0]])
    with torch.no_grad():
        pred = pipe({"view0": {"image": u0[None].float() / 255.0, "image_size": size},
                     "view1": {"image": u1[None].float() / 255.0, "image_size": size}})
    save("sg_pipeline", image0_u8=u0, image1_u8=u1,
         keypoints0=pred["keypoints0"], keypoints1=pred["keypoints1"],
         keypoint_scores0=pred["keypoint_scores0"], keypoint_scores1=pred["keypoint_scores1"],
         matches0=pred["matches0"], matches1=pred["matches1"],
         matching_scores0=pred["matching_scores0"], matching_scores1=pred["matching_scores1"],
         log_assignment_sub=pred["log_assignment"][:, ::4, ::4])


if __name__ == "__main__":
    main()
