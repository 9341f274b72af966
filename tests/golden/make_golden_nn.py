"""Generate golden vectors for the nearest-neighbour matcher by running the UNMODIFIED reference
(authoring container only; see make_golden.py for the import arrangement).

    python tests/golden/make_golden_nn.py
"""
import os
import sys
import warnings

REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path = [p for p in sys.path if os.path.abspath(p or ".") != REPO]
sys.path[:0] = ["/root/reference", os.path.join(REPO, "oracle", "_shim")]
sys.path.append(REPO)  # only for the conf stand-in's backend (no `gluefactory` shadowing: the reference is first)

import numpy as np  # noqa: E402
import torch  # noqa: E402

warnings.filterwarnings("ignore")
import gluefactory  # noqa: E402

assert gluefactory.__file__.startswith("/root/reference"), gluefactory.__file__
from gluefactory.models import get_model  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
torch.set_num_threads(8)


def descriptors(seed, b, m, n, d, noise):
    """Unit descriptors; image 1 holds noisy copies of a random subset of image 0 plus unrelated ones."""
    g = torch.Generator().manual_seed(seed)
    d0 = torch.nn.functional.normalize(torch.randn(b, m, d, generator=g), dim=-1)
    d1 = torch.nn.functional.normalize(torch.randn(b, n, d, generator=g), dim=-1)
    shared = min(m, n) * 2 // 3
    for i in range(b):
        src = torch.randperm(m, generator=g)[:shared]
        dst = torch.randperm(n, generator=g)[:shared]
        d1[i, dst] = torch.nn.functional.normalize(d0[i, src] + noise * torch.randn(shared, d, generator=g), dim=-1)
    return d0, d1


def main():
    cases = {
        "nn_a": dict(seed=11, b=2, m=160, n=190, d=128, noise=0.05, conf={}),
        "nn_b": dict(seed=12, b=2, m=190, n=130, d=128, noise=0.08,
                     conf={"ratio_thresh": 0.9, "distance_thresh": 0.9}),
        "nn_c": dict(seed=13, b=1, m=97, n=300, d=64, noise=0.08,
                     conf={"ratio_thresh": 0.95, "mutual_check": False}),
    }
    Model = get_model("matchers.nearest_neighbor_matcher")
    for name, c in cases.items():
        d0, d1 = descriptors(c["seed"], c["b"], c["m"], c["n"], c["d"], c["noise"])
        model = Model(c["conf"]).eval()
        with torch.no_grad():
            pred = model({"descriptors0": d0, "descriptors1": d1})
        arrays = {"descriptors0": d0.numpy(), "descriptors1": d1.numpy(),
                  "ratio_thresh": np.float64(c["conf"].get("ratio_thresh") or 0.0),
                  "distance_thresh": np.float64(c["conf"].get("distance_thresh") or 0.0),
                  "mutual_check": np.int64(c["conf"].get("mutual_check", True))}
        for k, v in pred.items():
            arrays["out_" + k] = v.numpy()
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **arrays)
        print(name, {k: getattr(v, "shape", None) for k, v in arrays.items()},
              "matched", int((pred["matches0"] > -1).sum()))


if __name__ == "__main__":
    main()
