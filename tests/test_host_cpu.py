"""CPU-side checks: the C-ABI library loads and exports every symbol include/kanvit_b200.h
declares (no compute calls), the modules mirror the reference's state_dict, argument validation
returns error codes instead of aborting."""
import ctypes as C
import json
import os
import re

import pytest
import torch

from conftest import GOLDEN, ROOT


def _lib():
    from kan_vit_b200 import _lib as L
    return L, L.load()


def test_library_exports_every_declared_symbol():
    L, lib = _lib()
    header = open(os.path.join(ROOT, "include", "kanvit_b200.h")).read()
    declared = set(re.findall(r"\b(kv_[a-z0-9_]+)\s*\(", header))
    declared -= {"kv_status", "kv_family"}
    assert declared, "no declarations found"
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    for name in L.EXPORTS:
        assert name in declared, f"{name} bound in Python but not declared in the header"
    assert lib.kv_version() == 1
    assert lib.kv_status_string(-2).decode() == "shape not supported"


def test_bad_arguments_return_codes_without_a_gpu():
    L, lib = _lib()
    d = L.KanDesc(L.KV_EFFKAN, 4, 8, 8, 1, 1, 9, 5, 3, 8, 8, 0, 0, 0, 0, 1.0, 8 * 12 + 4, 0)
    assert lib.kv_kan_fparam_count(C.byref(d)) == 8 * 12 + 4  # knots + the 4-float grid probe
    assert lib.kv_kan_fwd(C.byref(d), None, None, None, None, None, None, None, None, None, None, None) == -1
    d.M = 7  # inconsistent with grid_size + spline_order + 1
    assert lib.kv_kan_fwd(C.byref(d), 1, 1, None, 1, None, None, None, None, None, 1, None) == -2
    d.family = 42
    assert lib.kv_kan_fwd(C.byref(d), 1, 1, None, 1, None, None, None, None, None, 1, None) == -3
    assert lib.kv_patchify(1, 1, 2, 1, 28, 28, 5, None) == -2  # 28 % 5 != 0 (model.py:58-59)
    assert lib.kv_attn_fwd(1, None, 1, 1, 1, 50, 2, 4096, None) == -2
    assert lib.kv_cross_entropy(1, 1, 0, 10, -100, 1, 1, 1, None, None, None, None) == -1      # empty batch
    assert lib.kv_cross_entropy(None, 1, 4, 10, -100, 1, 1, 1, None, None, None, None) == -1   # null logits
    assert lib.kv_image_augment(None, None, None, None, 2, 32, 32, 3, 4, 1, 1, 1, None) == -1
    assert lib.kv_adam_step(16, 16, 16, 16, 6, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, 16, None) == -1   # n % 4
    assert lib.kv_adam_step(16, 20, 16, 16, 8, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, 16, None) == -4   # alignment
    assert lib.kv_adam_step(None, None, None, None, 0, 1e-3, 0.9, 0.999, 1e-8, 1.0, 0, None, None) == 0
    assert lib.kv_cross_entropy_workspace_floats() >= 2


@pytest.mark.parametrize("name,args", [
    ("cfg1_efficientkan", dict(chw=(1, 28, 28), n_patches=7, n_blocks=4, d_hidden=64, n_heads=2, out_d=10, type="efficientkan")),
    ("cfg3_fast", dict(chw=(3, 32, 32), n_patches=4, n_blocks=6, d_hidden=192, n_heads=8, out_d=10, type="fast")),
    ("cfg3_sine", dict(chw=(3, 32, 32), n_patches=4, n_blocks=6, d_hidden=192, n_heads=8, out_d=10, type="sine")),
    ("cfg2_cheby", dict(chw=(1, 28, 28), n_patches=7, n_blocks=4, d_hidden=64, n_heads=2, out_d=10, type="cheby")),
    ("vanilla", dict(chw=(1, 28, 28), n_patches=7, n_blocks=2, d_hidden=64, n_heads=2, out_d=10, type="vanilla")),
])
def test_state_dict_matches_reference(name, args):
    import kan_vit_b200 as K
    keys = json.load(open(os.path.join(GOLDEN, "state_dict_keys.json")))[name]
    sd = K.VisionTransformer(**args).state_dict()
    assert list(sd) == list(keys) or set(sd) == set(keys)
    for k, shape in keys.items():
        assert list(sd[k].shape) == shape, k


def test_same_seed_same_init_as_reference():
    """reset_parameters draws in the reference's order: with the reference's seed the golden
    weights are reproduced exactly (KANLinear lstsq fit included, within solver noise)."""
    import kan_vit_b200 as K
    from conftest import load_golden
    g = load_golden("layer_cheby")
    torch.manual_seed(2)
    m = K.ChebyKANLayer(16, 24, 4)
    assert torch.equal(m.cheby_coeffs.detach(), g.t("sd/cheby_coeffs"))
    g = load_golden("layer_sine")
    torch.manual_seed(4)
    m = K.SineKANLayer(16, 24)
    for k, v in m.state_dict().items():
        assert torch.equal(v, g.t("sd/" + k)), k
    g = load_golden("layer_effkan")
    torch.manual_seed(0)
    m = K.KANLinear(16, 24)
    assert torch.equal(m.base_weight.detach(), g.t("sd/base_weight"))
    assert torch.allclose(m.spline_weight.detach(), g.t("sd/spline_weight"), atol=1e-6)
    assert torch.equal(m.spline_scaler.detach(), g.t("sd/spline_scaler"))


def test_no_cpu_fallback():
    import kan_vit_b200 as K
    from kan_vit_b200._lib import KanVitError
    if torch.cuda.is_available():
        pytest.skip("checks the CPU-tensor error path")
    m = K.ChebyKANLayer(4, 4, 3)
    with pytest.raises(KanVitError):
        m(torch.zeros(2, 4))


def test_reference_quirks_kept():
    import kan_vit_b200 as K
    m = K.NaiveFourierKANLayer(4, 4, 3, smootorch_initialization=True)
    assert m.fouriercoeffs.shape == (2, 4, 4, 3)
    assert K.ChebyKANLayer(4, 6, 3).arange.dtype == torch.int64
    f = K.FastKANLayer(4, 6)
    assert not f.rbf.grid.requires_grad and "rbf.grid" in f.state_dict()
    s = K.SineKANLayer(4, 6, 'cpu', 5)
    assert s.freq.shape == (1, 1, 1, 5) and s.phase.shape == (1, 1, 4, 5) and "phase" in s.state_dict()
    v = K.VisionTransformer((1, 28, 28))
    assert "pos_embeddings" not in v.state_dict()
    K.MSA(16, 2, type="bogus")  # the reference swallows the error at construction (attention.py:174-176)


def test_tensor_core_coverage_table():
    """Which layers the bf16 mode serves with tensor-core kernels (kv_kan_tc_bytes is a pure host query):
    forward for M <= 24, weight gradient for every family, input gradient for M <= 16 and nn.Linear."""
    import ctypes as C
    from kan_vit_b200 import _lib as L
    from kan_vit_b200.ops import KanSpec
    lib = L.load()

    def sizes(spec, T=1000, flags=0):
        n_in = spec.n_slices * spec.in_features
        d = spec.desc(T, n_in, spec.n_groups * spec.out_features, extra_flags=L.KV_FLAG_BF16_TC | flags)
        a, b, c = C.c_size_t(0), C.c_size_t(0), C.c_size_t(0)
        rc = lib.kv_kan_tc_bytes(C.byref(d), C.byref(a), C.byref(b), C.byref(c))
        return rc, a.value, b.value, c.value

    dense = {
        "effkan": KanSpec(L.KV_EFFKAN, 768, 3072, grid_size=5, spline_order=3),
        "cheby": KanSpec(L.KV_CHEBY, 64, 64, grid_size=4),
        "fastkan": KanSpec(L.KV_FASTKAN, 192, 192, grid_size=8),
        "sine": KanSpec(L.KV_SINE, 192, 192, grid_size=5),
        "fourier": KanSpec(L.KV_FOURIER, 64, 64, grid_size=5),
        "linear": KanSpec(L.KV_LINEAR, 768, 3072),
    }
    for name, spec in dense.items():
        rc, fwd, bwd_in, dyimg = sizes(spec, flags=L.KV_FLAG_UNIFORM_GRID if name == "effkan" else 0)
        assert rc == 0 and fwd > 0 and bwd_in > 0 and dyimg > 0, name
    # any number of basis functions (<= KV_MAX_M) runs on the tensor cores: the forward and the weight gradient walk the
    # bases in parts, the input gradient takes two accumulator passes above 24 and two launches (basis splits) above 32
    for G in (20, 28, 40, 64):
        rc, fwd, bwd_in, dyimg = sizes(KanSpec(L.KV_SINE, 192, 192, grid_size=G))
        assert rc == 0 and fwd > 0 and bwd_in > 0 and dyimg > 0, G
    # per-head q/k/v maps: merged (effkan on a uniform grid) or served as groups sharing a slice (fastkan)
    for spec, fl in ((KanSpec(L.KV_EFFKAN, 64, 64, n_groups=36, x_share=3, grid_size=5, spline_order=3),
                      L.KV_FLAG_UNIFORM_GRID),
                     (KanSpec(L.KV_FASTKAN, 24, 24, n_groups=24, x_share=3, grid_size=8), 0)):
        rc, fwd, bwd_in, dyimg = sizes(spec, flags=fl)
        assert rc == 0 and fwd > 0 and bwd_in > 0 and dyimg > 0


def test_metrics_from_confusion_match_sklearn():
    """utils.py:38-40: accuracy / balanced accuracy / weighted F1 are recomputed from a confusion matrix."""
    import numpy as np
    from sklearn.metrics import accuracy_score, balanced_accuracy_score, f1_score
    from kan_vit_b200.metrics import metrics_from_confusion
    rng = np.random.default_rng(0)
    for n_classes, n in ((10, 500), (100, 3000), (5, 40)):
        y = rng.integers(0, n_classes, n)
        y[y == 3] = 2                                    # a class without samples that is still predicted
        p = np.where(rng.random(n) < 0.6, y, rng.integers(0, n_classes, n))
        conf = np.zeros((n_classes, n_classes), np.int64)
        np.add.at(conf, (y, p), 1)
        acc, bal, f1 = metrics_from_confusion(conf)
        assert abs(acc - accuracy_score(y, p)) < 1e-12
        assert abs(bal - balanced_accuracy_score(y, p)) < 1e-12
        assert abs(f1 - f1_score(y, p, average="weighted")) < 1e-12


def test_kanlinear_maintenance_ops_match_reference():
    """update_grid / curve2coeff / regularization_loss of the drop-in KANLinear (plain torch, device-agnostic)
    against values recorded from the reference (models/effkan.py:134-264)."""
    from conftest import load_golden, relerr
    import kan_vit_b200 as K
    g = load_golden("maint_effkan")
    m = K.KANLinear(6, 5)
    m.load_state_dict(g.sd())
    assert abs(float(m.regularization_loss(1.0, 1.0)) - float(g.raw["reg_before"])) < 1e-5 * abs(float(g.raw["reg_before"]))
    m.update_grid(g.t("x"))
    assert relerr(m.grid, g.t("grid_after")) < 1e-6
    assert relerr(m.spline_weight, g.t("spline_weight_after")) < 1e-3      # lstsq: conditioning of the refit
    assert abs(float(m.regularization_loss(0.5, 2.0)) - float(g.raw["reg_after"])) < 1e-4 * abs(float(g.raw["reg_after"]))
    assert not m.grid_is_uniform()
