"""Training losses are outside the forward-only scope of this build.

`NLLLoss` keeps the constructor signature the reference uses
(reference: gluefactory/models/utils/losses.py:29-37) so that `LightGlue.__init__`
builds; calling it raises.
"""
from torch import nn

from glue_factory_b200._conf import get_conf_backend

_, OmegaConf = get_conf_backend()


class NLLLoss(nn.Module):
    default_conf = {"nll_balancing": 0.5, "gamma_f": 0.0}

    def __init__(self, conf):
        super().__init__()
        self.conf = OmegaConf.merge(self.default_conf, conf)

    def forward(self, pred, data, weights=None):
        raise NotImplementedError("training losses are out of scope for the B200 forward path")
