"""Golden vectors for the rows either side of the path, produced by the UNMODIFIED
reference (authoring container only; see make_golden.py for the import arrangement):

  prep_*.npz    gluefactory.utils.image.ImagePreprocessor on crops of the boat image
  export_*.npz  gluefactory.utils.export_predictions.export_predictions on recorded predictions
  eval_*.npz    gluefactory.eval.utils.eval_matches_homography / eval_matches_epipolar / eval_homography_dlt

kornia and h5py are absent: oracle/_shim/kornia routes the two kornia calls (`geometry.transform.resize`,
`geometry.homography.find_homography_dlt`) to the oracle's restatement of kornia's published algorithms, and
oracle/_shim/h5py keeps the written groups in memory.

    python tests/golden/make_golden_io.py
"""
import os
import sys
import warnings

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != REPO]
sys.path[:0] = ["/root/reference", os.path.join(REPO, "oracle", "_shim")]
sys.path.append(REPO)  # `oracle` (for the shims) and the conf stand-in's backend; the reference stays first

import numpy as np  # noqa: E402
import torch  # noqa: E402

warnings.filterwarnings("ignore")
import gluefactory  # noqa: E402

assert gluefactory.__file__.startswith("/root/reference"), gluefactory.__file__
import h5py  # noqa: E402  (the in-memory stand-in)
from gluefactory.eval import utils as eval_utils  # noqa: E402
from gluefactory.geometry.wrappers import Camera, Pose  # noqa: E402
from gluefactory.utils.export_predictions import export_predictions  # noqa: E402
from gluefactory.utils.image import ImagePreprocessor, read_image, numpy_image_to_torch  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
torch.set_num_threads(8)


# ---- pre-processing ------------------------------------------------------------------------------------
def make_prep():
    rgb = read_image(os.path.join(OUT, "assets", "boat1.png"))          # 680 x 850 x 3 uint8
    gray = read_image(os.path.join(OUT, "assets", "boat2.png"), grayscale=True)
    cases = {
        "prep_a": (rgb[100:340, 200:520].copy(), dict(resize=96, side="short")),                  # 240x320 -> 96x128, blur
        "prep_b": (gray[50:350, 30:250].copy(), dict(resize=128, side="long", square_pad=True, add_padding_mask=True)),
        "prep_c": (rgb[0:90, 0:150].copy(), dict(resize=200, side="long")),                       # up-scaling: no blur
        "prep_d": (gray[10:210, 400:700].copy(), dict(resize=[75, 170])),                         # explicit (h, w)
        "prep_e": (rgb[300:500, 100:421].copy(), dict(resize=150, side="long", edge_divisible_by=8, antialias=False)),
        "prep_f": (gray[0:64, 0:96].copy(), dict(resize=None, square_pad=True)),                  # conversion + padding only
        "prep_g": (rgb[0:330, 0:440].copy(), dict(resize=60, side="vert")),                       # factor 5.5: 11 taps
    }
    for name, (u8, conf) in cases.items():
        out = ImagePreprocessor(conf)(numpy_image_to_torch(u8))
        rec = {"input_u8": u8, "conf_keys": np.array(list(conf.keys())),
               "conf_vals": np.array([repr(v) for v in conf.values()]),
               "image": out["image"].numpy(), "scales": out["scales"].numpy(), "image_size": out["image_size"],
               "transform": np.asarray(out["transform"], dtype=np.float64),
               "original_image_size": out["original_image_size"]}
        if "padding_mask" in out:
            rec["padding_mask"] = out["padding_mask"].numpy()
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
        print(name, u8.shape, "->", out["image"].shape, out["scales"].tolist())


# ---- export --------------------------------------------------------------------------------------------
class _Replay(torch.nn.Module):
    def __init__(self, preds):
        super().__init__()
        self.preds, self.i = preds, 0

    def forward(self, data):
        self.i += 1
        return {k: v.clone() for k, v in self.preds[self.i - 1].items()}


def make_export():
    g = torch.Generator().manual_seed(11)
    K0, K1 = 37, 29
    keys = ["keypoints0", "keypoints1", "keypoint_scores0", "keypoint_scores1", "matches0", "matches1",
            "matching_scores0", "matching_scores1"]
    loader, preds = [], []
    for i in range(3):
        pred = {"keypoints0": torch.rand(1, K0, 2, generator=g) * 640, "keypoints1": torch.rand(1, K1, 2, generator=g) * 480,
                "keypoint_scores0": torch.rand(1, K0, generator=g), "keypoint_scores1": torch.rand(1, K1, generator=g),
                "matches0": torch.randint(-1, K1, (1, K0), generator=g), "matches1": torch.randint(-1, K0, (1, K1), generator=g),
                "matching_scores0": torch.rand(1, K0, generator=g), "matching_scores1": torch.rand(1, K1, generator=g),
                "log_assignment": torch.randn(1, K0 + 1, K1 + 1, generator=g)}
        data = {"view0": {"scales": torch.rand(1, 2, generator=g) * 0.6 + 0.3},
                "view1": {"scales": torch.rand(1, 2, generator=g) * 0.6 + 0.3}, "name": [f"pair{i}/a_b"]}
        loader.append(data)
        preds.append(pred)
    for tag, half in (("export_f32", False), ("export_f16", True)):
        export_predictions(loader, _Replay(preds), os.path.join("/tmp", tag + ".h5"), as_half=half, keys=keys)
        written = h5py.LAST_FILE
        rec = {}
        for i, (data, pred) in enumerate(zip(loader, preds)):
            for k, v in pred.items():
                rec[f"in{i}/{k}"] = v.numpy()
            rec[f"in{i}/scales0"] = data["view0"]["scales"].numpy()
            rec[f"in{i}/scales1"] = data["view1"]["scales"].numpy()
            for k, v in written[data["name"][0]].items():
                rec[f"out{i}/{k}"] = v
        rec["names"] = np.array([d["name"][0] for d in loader])
        rec["keys"] = np.array(keys)
        np.savez_compressed(os.path.join(OUT, tag + ".npz"), **rec)
        print(tag, sorted(written.keys()), {k: v.dtype for k, v in written[loader[0]["name"][0]].items()})


# ---- match metrics -------------------------------------------------------------------------------------
def _pair(g, M, N, n_match, warp, noise, outliers):
    """keypoints0 random; `n_match` of them matched to warp(kp0) + noise, a share of them to random points."""
    kp0 = torch.rand(M, 2, generator=g) * torch.tensor([640.0, 480.0])
    kp1 = torch.rand(N, 2, generator=g) * torch.tensor([640.0, 480.0])
    m0 = torch.full((M,), -1, dtype=torch.int64)
    src = torch.randperm(M, generator=g)[:n_match]
    dst = torch.randperm(N, generator=g)[:n_match]
    good = torch.rand(n_match, generator=g) > outliers
    kp1[dst[good]] = warp(kp0[src[good]]) + noise * torch.randn(int(good.sum()), 2, generator=g)
    m0[src] = dst
    ms0 = torch.where(m0 > -1, torch.rand(M, generator=g) * 0.9 + 0.1, torch.zeros(M))
    return kp0, kp1, m0, ms0


def make_eval():
    g = torch.Generator().manual_seed(23)
    rec = {}
    # homographies
    Hs = [torch.tensor([[1.02, 0.03, 5.0], [-0.02, 0.98, -3.0], [1e-5, -2e-5, 1.0]]),
          torch.tensor([[0.9, -0.2, 40.0], [0.15, 1.1, 10.0], [1e-4, 5e-5, 1.0]]),
          torch.tensor([[1.0, 0.0, 0.0], [0.0, 1.0, 0.0], [0.0, 0.0, 1.0]]),
          torch.tensor([[0.9, -0.2, 40.0], [0.15, 1.1, 10.0], [1e-4, 5e-5, 1.0]]),      # no outliers: well-posed DLT
          torch.tensor([[1.05, 0.1, -20.0], [-0.08, 0.97, 15.0], [-3e-5, 2e-5, 1.0]])]
    for i, (H, M, N, n, noise, outl) in enumerate(zip(Hs, (300, 512, 64, 400, 2048), (280, 400, 64, 420, 2048),
                                                      (150, 300, 3, 250, 1500), (0.8, 2.0, 0.1, 0.5, 1.0),
                                                      (0.2, 0.4, 0.0, 0.0, 0.0))):
        def warp(p, H=H):
            q = torch.cat([p, torch.ones(len(p), 1)], -1) @ H.T
            return q[:, :2] / q[:, 2:]

        kp0, kp1, m0, ms0 = _pair(g, M, N, n, warp, noise, outl)
        data = {"H_0to1": H, "view0": {"image_size": torch.tensor([640.0, 480.0])}}
        pred = {"keypoints0": kp0, "keypoints1": kp1, "matches0": m0, "matching_scores0": ms0}
        res = eval_utils.eval_matches_homography(data, pred)
        res.update(eval_utils.eval_homography_dlt(data, pred))
        for k, v in {**pred, "H_0to1": H}.items():
            rec[f"h{i}/{k}"] = v.numpy()
        for k, v in res.items():
            rec[f"h{i}/res/{k}"] = np.float64(v)
        print("homography", i, res)
    rec["h_count"] = np.int64(len(Hs))
    # relative poses, pinhole cameras
    for i, (M, N, n, noise, outl) in enumerate(((400, 380, 200, 0.3, 0.3), (256, 256, 120, 1.5, 0.1))):
        ang = 0.1 + 0.2 * i
        R = torch.tensor([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]], dtype=torch.float32)
        t = torch.tensor([0.5, -0.1, 0.2 + i])
        cam0 = torch.tensor([640.0, 480.0, 500.0 + 20 * i, 510.0, 320.0, 240.0])
        cam1 = torch.tensor([640.0, 480.0, 480.0, 470.0 - 10 * i, 330.0, 235.0])
        depth = torch.rand(M, generator=g) * 4 + 2

        def warp(p, idx=[0]):
            # consistent two-view geometry: back-project with a per-call depth, transform, project
            d = depth[: len(p)]
            x = torch.cat([(p - cam0[4:6]) / cam0[2:4], torch.ones(len(p), 1)], -1) * d[:, None]
            y = x @ R.T + t
            return y[:, :2] / y[:, 2:] * cam1[2:4] + cam1[4:6]

        kp0, kp1, m0, ms0 = _pair(g, M, N, n, warp, noise, outl)
        data = {"view0": {"camera": Camera(cam0[None])}, "view1": {"camera": Camera(cam1[None])},
                "T_0to1": Pose.from_Rt(R[None], t[None])}
        pred = {"keypoints0": kp0, "keypoints1": kp1, "matches0": m0, "matching_scores0": ms0}
        res = eval_utils.eval_matches_epipolar(data, pred)
        for k, v in {**pred, "R": R, "t": t, "cam0": cam0, "cam1": cam1}.items():
            rec[f"e{i}/{k}"] = v.numpy()
        for k, v in res.items():
            rec[f"e{i}/res/{k}"] = np.float64(float(v))
        print("epipolar", i, {k: float(v) for k, v in res.items()})
    rec["e_count"] = np.int64(2)
    np.savez_compressed(os.path.join(OUT, "eval_metrics.npz"), **rec)


if __name__ == "__main__":
    make_prep()
    make_export()
    make_eval()
