"""CPU tests of the host-side boundary: conf subset, registry, BaseModel contract, state-dict schemas."""
import pytest
import torch

from glue_factory_b200 import synth
from glue_factory_b200._conf import ConfigAttributeError, OmegaConf, ReadonlyConfigError


def test_conf_merge_struct_readonly():
    base = OmegaConf.create({"a": 1, "nested": {"x": 1, "y": [1, 2]}})
    merged = OmegaConf.merge(base, {"nested": {"x": 5}, "extra": True})
    assert merged.nested.x == 5 and merged.nested.y == [1, 2] and merged.extra is True and base.nested.x == 1
    OmegaConf.set_struct(merged, True)
    OmegaConf.set_readonly(merged, True)
    with pytest.raises(ConfigAttributeError):
        merged.missing
    with pytest.raises(ReadonlyConfigError):
        merged.a = 2
    assert OmegaConf.to_container(merged)["nested"] == {"x": 5, "y": [1, 2]}
    assert OmegaConf.from_dotlist(["model.matcher.name=foo", "k=3"]).model.matcher.name == "foo"


def test_registry_resolves_like_the_reference():
    from gluefactory.models import get_model
    from gluefactory.models.base_model import BaseModel

    sp = get_model("extractors.superpoint_open")
    lg = get_model("matchers.lightglue")
    assert issubclass(sp, BaseModel) and sp.__module__ == "glue_factory_b200.models.superpoint_open"
    assert lg.__name__ == "LightGlue"
    assert get_model("glue_factory_b200.models.superpoint_open") is sp       # swap-in route (ii)
    assert get_model("glue_factory_b200.models.lightglue") is lg
    assert get_model("two_view_pipeline").__name__ == "TwoViewPipeline"
    with pytest.raises(RuntimeError):
        get_model("no_such_model")


def test_state_dict_schemas_and_conf_defaults(tmp_path, sp_weights, lg_weights):
    from gluefactory.models import get_model

    torch.save(sp_weights, tmp_path / "sp.pth")
    torch.save(lg_weights, tmp_path / "lg.pth")
    sp = get_model("extractors.superpoint_open")({"weights": str(tmp_path / "sp.pth"), "unknown_key": 1})
    assert list(sp.state_dict().keys()) == list(sp_weights.keys())            # 84 tensors, reference order
    assert sp.conf.nms_radius == 4 and sp.conf.detection_threshold == 0.005 and sp.is_initialized()
    lg = get_model("matchers.lightglue")({"weights": str(tmp_path / "lg.pth")})
    assert set(lg.state_dict().keys()) == set(lg_weights.keys())
    assert lg.conf.n_layers == 9 and lg.conf.filter_threshold == 0.0 and lg.conf.loss.nll_balancing == 0.5
    for k, v in lg.state_dict().items():
        assert torch.equal(v, lg_weights[k]), k
    with pytest.raises(AssertionError):
        lg({"keypoints0": torch.zeros(1, 4, 2)})                               # missing keys
    with pytest.raises(NotImplementedError):
        sp.loss({}, {})
    # legacy checkpoint key names are renamed (lightglue.py:374-380)
    legacy = {k.replace("transformers.0.self_attn", "self_attn.0"): v for k, v in lg_weights.items()}
    torch.save(legacy, tmp_path / "legacy.pth")
    lg2 = get_model("matchers.lightglue")({"weights": str(tmp_path / "legacy.pth")})
    assert torch.equal(lg2.transformers[0].self_attn.Wqkv.weight, lg_weights["transformers.0.self_attn.Wqkv.weight"])


def test_pipeline_requires_views(tmp_path, sp_weights):
    from gluefactory.models import get_model

    torch.save(sp_weights, tmp_path / "sp.pth")
    pipe = get_model("two_view_pipeline")({"extractor": {"name": "extractors.superpoint_open",
                                                         "weights": str(tmp_path / "sp.pth")}})
    with pytest.raises(AssertionError):
        pipe({"view0": {}})


def test_qkv_repacking_matches_reference_layout(lg_weights):
    """The host de-interleaves Wqkv from (head, dim, {q,k,v}) to [q | k | v] (lightglue.py:151)."""
    w = lg_weights["transformers.0.self_attn.Wqkv.weight"]
    x = torch.randn(5, 256)
    qkv = (x @ w.t()).view(5, 4, 64, 3)
    perm = torch.arange(768).view(4, 64, 3).permute(2, 0, 1).reshape(-1)
    packed = x @ w[perm].t()
    assert torch.allclose(packed[:, :256].view(5, 4, 64), qkv[..., 0])
    assert torch.allclose(packed[:, 512:].view(5, 4, 64), qkv[..., 2])


def test_synthetic_pairs_are_deterministic():
    a = synth.synthetic_pair(5, 64, 80)
    b = synth.synthetic_pair(5, 64, 80)
    assert all(torch.equal(x, y) for x, y in zip(a, b)) and a[0].shape == (1, 64, 80)
    assert 0.0 <= a[1].min() and a[1].max() <= 1.0
