"""Matcher metrics are evaluated by the reference's eval harness, not by this forward path
(reference: gluefactory/models/utils/metrics.py)."""


def matcher_metrics(pred, data, prefix="", prefix_gt=None):
    raise NotImplementedError("matcher metrics are out of scope for the B200 forward path")
