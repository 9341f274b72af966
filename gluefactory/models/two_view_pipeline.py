"""Two-view pipeline: extractor on each view, then matcher / filter / solver
(reference: gluefactory/models/two_view_pipeline.py:21-114).

Matches convention: ``matches0[i]`` is the index in image 1 matched to keypoint i of image 0,
-1 when unmatched.
"""
from glue_factory_b200._conf import get_conf_backend

from . import get_model
from .base_model import BaseModel

_, OmegaConf = get_conf_backend()
to_ctr = OmegaConf.to_container


class TwoViewPipeline(BaseModel):
    default_conf = {
        "extractor": {"name": None, "trainable": False},
        "matcher": {"name": None},
        "filter": {"name": None},
        "solver": {"name": None},
        "ground_truth": {"name": None},
        "allow_no_extract": False,
        "run_gt_in_forward": False,
    }
    required_data_keys = ["view0", "view1"]
    strict_conf = False
    components = ["extractor", "matcher", "filter", "solver", "ground_truth"]

    def _init(self, conf):
        for part in self.components:
            if conf[part].name:
                setattr(self, part, get_model(conf[part].name)(to_ctr(conf[part])))

    def extract_view(self, data, i):
        view = data[f"view{i}"]
        pred = view.get("cache", {})
        cached = len(pred) > 0 and self.conf.allow_no_extract
        if self.conf.extractor.name and not cached:
            pred = {**pred, **self.extractor(view)}
        elif self.conf.extractor.name and not self.conf.allow_no_extract:
            pred = {**pred, **self.extractor({**view, **pred})}
        return pred

    def _forward(self, data):
        pred = {}
        for i in ("0", "1"):
            pred.update({k + i: v for k, v in self.extract_view(data, i).items()})
        for part in ("matcher", "filter", "solver"):
            if self.conf[part].name:
                pred = {**pred, **getattr(self, part)({**data, **pred})}
        if self.conf.ground_truth.name and self.conf.run_gt_in_forward:
            gt = self.ground_truth({**data, **pred})
            pred.update({f"gt_{k}": v for k, v in gt.items()})
        return pred

    def loss(self, pred, data):
        losses, metrics, total = {}, {}, 0
        if self.conf.ground_truth.name and not self.conf.run_gt_in_forward:
            gt = self.ground_truth({**data, **pred})
            pred.update({f"gt_{k}": v for k, v in gt.items()})
        for part in self.components:
            enabled = self.conf[part].get("apply_loss", True) if "apply_loss" in self.conf[part].keys() else True
            if not (self.conf[part].name and enabled):
                continue
            try:
                part_losses, part_metrics = getattr(self, part).loss(pred, {**pred, **data})
            except NotImplementedError:
                continue
            losses.update(part_losses)
            metrics.update(part_metrics)
            total = part_losses["total"] + total
        return {**losses, "total": total}, metrics
