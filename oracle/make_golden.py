"""Generate tests/golden/*.npz by running the UNMODIFIED reference.  TEST INFRASTRUCTURE.

Run in the build container (where /root/reference exists):

    python oracle/make_golden.py

The reference is imported from /root/reference (never copied); every case is
seeded, so re-running reproduces the committed files.  Two labelled shims, both
in this harness only:
  * cheby ViT: the reference crashes at model.py:150 because ChebyKANLayer
    returns 2-D; the harness wraps linear_mapper so its output is reshaped to
    (B, P^2, d).
  * fourier: model.py:74 passes a wrong kwarg; fourier is layer-level only.
The KAN-MLP extension case replaces ``ff`` of a reference TransformerBlock by
two reference KANLinear modules.
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def _np(t):
    return t.detach().cpu().numpy()


def save_case(name, module, x, extra=None, out_fn=None, f64=True):
    """Run module fwd+bwd on x with a seeded upstream grad; dump everything."""
    x = x.clone().requires_grad_(True)
    y = module(x) if out_fn is None else out_fn(module, x)
    g = torch.Generator().manual_seed(1234)
    dy = torch.randn(y.shape, generator=g)
    module.zero_grad()
    y.backward(dy)
    rec = {"x": _np(x), "dy": _np(dy), "y": _np(y), "gx": _np(x.grad)}
    for k, v in module.state_dict().items():
        rec["sd/" + k] = _np(v)
    for k, p in module.named_parameters():
        if p.grad is not None:
            rec["g/" + k] = _np(p.grad)
    # float64 ground truth of the forward (settles summation-order disputes)
    if f64:  # (the reference's FlashAttentionFunction builds fp32 state tensors and cannot run in float64)
        m64 = module.double()
        with torch.no_grad():
            y64 = m64(x.detach().double()) if out_fn is None else out_fn(m64, x.detach().double())
        rec["y64"] = _np(y64)
        module.float()
    if extra:
        rec.update(extra)
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print(f"{name}: y{tuple(y.shape)} {sum(v.nbytes for v in rec.values())/1024:.1f} KiB")


def flash_cases():
    """FlashAttention module / type="flash-attn" (SURVEY.md section 8f rank 2):
    python oracle/make_golden.py --flash-only"""
    sys.path.insert(0, REF)
    import model as ref_model  # noqa
    from attention import FlashAttention  # noqa

    def xin(shape, seed, scale=1.0):
        return torch.randn(shape, generator=torch.Generator().manual_seed(seed)) * scale

    def bmask(shape, seed, p=0.7):
        return torch.rand(shape, generator=torch.Generator().manual_seed(seed)) < p

    def case(name, seed, x, causal=False, context=None, mask=None, qb=None, kb=None):
        torch.manual_seed(seed)
        m = FlashAttention(dim=16, heads=2, dim_head=8, causal=causal)
        extra = {"causal": np.int64(causal)}
        if context is not None:
            extra["context"] = _np(context)
        if mask is not None:
            extra["mask"] = _np(mask)
        save_case(name, m, x, extra=extra, f64=False,
                  out_fn=lambda mod, xx: mod(xx, context=None if context is None else context.to(xx.dtype),
                                             mask=mask, q_bucket_size=qb, k_bucket_size=kb))

    case("flash_self", 60, xin((3, 5, 16), 61))
    km = bmask((3, 5), 62)
    km[1] = False                                   # one sample with every key hidden: o = 0
    case("flash_keymask", 63, xin((3, 5, 16), 64), mask=km)
    case("flash_causal", 65, xin((3, 6, 16), 66), causal=True)
    # more keys than queries + causal: the first Nk-Nq query rows see no key and attend uniformly
    case("flash_cross_causal", 67, xin((2, 4, 16), 68), causal=True, context=xin((2, 7, 16), 69))
    case("flash_cross_mask", 70, xin((2, 6, 16), 71), context=xin((2, 4, 16), 72), mask=bmask((2, 4), 73))
    case("flash_fullmask_causal_buckets", 74, xin((2, 7, 16), 75), causal=True, mask=bmask((2, 1, 7, 7), 76),
         qb=3, kb=2)

    torch.manual_seed(40)
    vit = ref_model.VisionTransformer((1, 8, 8), n_patches=2, n_blocks=2, d_hidden=16, n_heads=2, out_d=10,
                                      type="flash-attn")
    x = xin((3, 1, 8, 8), 41).clone().requires_grad_(True)
    y = vit(x)
    dy = torch.randn(y.shape, generator=torch.Generator().manual_seed(1234))
    y.backward(dy)
    rec = {"x": _np(x), "dy": _np(dy), "y": _np(y), "gx": _np(x.grad), "pos_embeddings": _np(vit.pos_embeddings)}
    for k, v in vit.state_dict().items():
        rec["sd/" + k] = _np(v)
    for k, p in vit.named_parameters():
        if p.grad is not None:
            rec["g/" + k] = _np(p.grad)
    np.savez_compressed(os.path.join(OUT, "vit_flash-attn.npz"), **rec)
    print(f"vit_flash-attn: y{tuple(y.shape)}")


def maintenance_case():
    """KANLinear.update_grid / curve2coeff / regularization_loss (models/effkan.py:134-264):
    python oracle/make_golden.py --maint-only"""
    sys.path.insert(0, REF)
    from models.effkan import KANLinear  # noqa
    torch.manual_seed(80)
    m = KANLinear(6, 5)
    g = torch.Generator().manual_seed(81)
    x = torch.randn(64, 6, generator=g) * torch.tensor([0.3, 0.6, 1.0, 1.5, 0.8, 0.2]) + torch.tensor([0., .2, -.3, 0., .5, 0.])
    probe = torch.randn(16, 6, generator=g) * 0.5
    rec = {"x": _np(x), "probe": _np(probe)}
    for k, v in m.state_dict().items():
        rec["sd/" + k] = _np(v).copy()      # update_grid below writes these tensors in place
    rec["reg_before"] = _np(m.regularization_loss(1.0, 1.0))
    rec["y_before"] = _np(m(probe))
    m.update_grid(x)
    rec["grid_after"] = _np(m.grid)
    rec["spline_weight_after"] = _np(m.spline_weight)
    rec["reg_after"] = _np(m.regularization_loss(0.5, 2.0))
    rec["y_after"] = _np(m(probe))
    rec["y_after_x"] = _np(m(x))
    np.savez_compressed(os.path.join(OUT, "maint_effkan.npz"), **rec)
    print("maint_effkan:", {k: v.shape for k, v in rec.items() if not k.startswith("sd/")})


def big_cases():
    """Parity fixtures at BASELINE config 3 / 4 geometry: python oracle/make_golden.py --big

    Weights and inputs come from tests/seeded.py (regenerated from the seed by the tests, never stored); the
    UNMODIFIED reference modules run on them and a sub-sample of every output / gradient is committed
    (about 4096 values per tensor at an odd stride -- 384 per parameter gradient of the multi-layer models -- plus the tensor's max |value| for the norm-wise error)."""
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
    import model as ref_model  # noqa
    from models.effkan import KANLinear  # noqa
    from seeded import seeded_input, seeded_state, subsample  # noqa

    def run(module, x, dy):
        x = x.clone().requires_grad_(True)
        y = module(x)
        module.zero_grad()
        y.backward(dy.to(y.dtype))
        out = {"y": y.detach(), "gx": x.grad}
        for k, p in module.named_parameters():
            if p.grad is not None:
                out["g/" + k] = p.grad.detach().clone()
        return out

    def record(name, module, x, seed):
        """Stored per tensor: the float64 run of the reference (ground truth, sub-sampled), its max |value|, and
        `@ref32err` = the norm-wise distance of the reference's own float32 run from it (how well-conditioned the
        tensor is in fp32: the SineKAN ViT's input gradient is 1e-3 away from its own fp64 value)."""
        module.load_state_dict(seeded_state(module.state_dict(), seed))
        with torch.no_grad():
            shape = module(x).shape
        dy = seeded_input(tuple(shape), seed + 7)
        r32 = run(module, x, dy)
        module.double()
        if hasattr(module, "patchify"):
            # labelled shim, float64 ground-truth run only: patchify allocates float32 whatever the input dtype
            # (model.py:116), which the float64 LayerNorm / Linear behind it reject
            orig = module.patchify
            module.patchify = lambda images, n: orig(images, n).double()
        r64 = run(module, x.double(), dy)
        if hasattr(module, "patchify"):
            del module.patchify
        module.float()
        rec = {}
        n_par = len(list(module.parameters()))
        for key, t in r64.items():
            v, k = subsample(t, 4096 if (n_par < 8 or not key.startswith("g/")) else 384)
            rec[key] = _np(v).astype(np.float32)   # the float64 values, stored to float32 precision
            amax = t.abs().max().item()
            e32 = (r32[key].double() - t).abs().max().item() / (amax if amax > 0 else 1.0)
            rec[key + "@meta"] = np.array([k, amax, e32], dtype=np.float64)   # stride, max |value|, ref32err
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
        worst = max((float(rec[k][2]), k) for k in rec if k.endswith("@meta") and float(rec[k][1]) > 1e-9)
        print(f"{name}: y{tuple(shape)} {sum(v.nbytes for v in rec.values())/1024:.1f} KiB; reference fp32 vs fp64: "
              f"y {rec['y@meta'][2]:.1e}, gx {rec['gx@meta'][2]:.1e}, worst {worst[0]:.1e} {worst[1]}")

    torch.manual_seed(0)
    record("big_effkan_768x3072_T300", KANLinear(768, 3072), seeded_input((300, 768), 101), 100)
    record("big_effkan_3072x768_T300", KANLinear(3072, 768), seeded_input((300, 3072), 111), 110)
    blk = ref_model.TransformerBlock(768, 12, feedforward_dim=3072, attn_type="efficientkan")
    blk.ff = torch.nn.Sequential(KANLinear(768, 3072), KANLinear(3072, 768))  # KAN-MLP extension
    record("big_block_cfg4_kanmlp_B2", blk, seeded_input((2, 197, 768), 121), 120)
    for kind in ("fast", "sine"):
        vit = ref_model.VisionTransformer((3, 32, 32), n_patches=4, n_blocks=6, d_hidden=192, n_heads=8, out_d=10,
                                          type=kind)
        record(f"big_vit_cfg3_{kind}_B4", vit, seeded_input((4, 3, 32, 32), 131), 130)
    vit = ref_model.VisionTransformer((3, 224, 224), n_patches=14, n_blocks=2, d_hidden=768, n_heads=12, out_d=10,
                                      type="efficientkan")
    for b in vit.blocks:
        b.ff = torch.nn.Sequential(KANLinear(768, 3072), KANLinear(3072, 768))
    record("big_vit_cfg4_2blk_kanmlp_B2", vit, seeded_input((2, 3, 224, 224), 141), 140)


def main():
    if "--big" in sys.argv:
        big_cases()
        return
    if "--flash-only" in sys.argv:
        flash_cases()
        return
    if "--maint-only" in sys.argv:
        maintenance_case()
        return
    sys.path.insert(0, REF)
    import model as ref_model  # noqa
    from attention import MSA  # noqa
    from models.cheby import ChebyKANLayer  # noqa
    from models.effkan import KANLinear  # noqa
    from models.fastkan import FastKANLayer  # noqa
    from models.nfkan import NaiveFourierKANLayer  # noqa
    from models.sinekan import SineKANLayer  # noqa

    # ---------------- layers ----------------
    def xin(shape, seed, scale=1.0):
        return torch.randn(shape, generator=torch.Generator().manual_seed(seed)) * scale

    torch.manual_seed(0)
    m = KANLinear(16, 24)
    x = xin((3, 5, 16), 1, 1.3)
    # exact knots, out-of-span values and the span edges (effkan.py:115 half-open test)
    x[0, 0, :8] = torch.tensor([-2.2, -1.0, -0.2, 0.6, 2.2, 2.5, -3.0, 1.0])
    save_case("layer_effkan", m, x)

    torch.manual_seed(1)
    m = KANLinear(8, 12, grid_size=10, spline_order=2, enable_standalone_scale_spline=False)
    with torch.no_grad():  # non-uniform, per-feature knots (what update_grid would leave behind)
        jitter = torch.rand(m.grid.shape, generator=torch.Generator().manual_seed(5)) * 0.05
        m.grid.add_(jitter.cumsum(1))
    save_case("layer_effkan_g10k2_nonuniform", m, xin((7, 8), 2),
              extra={"spline_order": np.int64(2)})

    torch.manual_seed(2)
    save_case("layer_cheby", ChebyKANLayer(16, 24, 4), xin((3, 5, 16), 3, 1.5))
    torch.manual_seed(12)
    save_case("layer_cheby_deg7", ChebyKANLayer(8, 8, 7), xin((9, 8), 13))

    torch.manual_seed(3)
    m = FastKANLayer(16, 24)
    with torch.no_grad():
        m.layernorm.weight.add_(0.1 * torch.randn(16))
        m.layernorm.bias.add_(0.1 * torch.randn(16))
    save_case("layer_fastkan", m, xin((3, 5, 16), 4))

    torch.manual_seed(4)
    save_case("layer_sine", SineKANLayer(16, 24), xin((3, 5, 16), 5))
    torch.manual_seed(5)
    save_case("layer_sine_g4", SineKANLayer(8, 8, grid_size=4), xin((6, 8), 6))

    torch.manual_seed(6)
    save_case("layer_fourier", NaiveFourierKANLayer(16, 24, 5), xin((3, 5, 16), 7))

    # ---------------- MSA / block ----------------
    for kind in ("vanilla", "efficientkan", "cheby", "fast", "sine"):
        torch.manual_seed(20)
        save_case(f"msa_{kind}", MSA(16, n_heads=2, type=kind), xin((3, 5, 16), 21))

    torch.manual_seed(30)
    blk = ref_model.TransformerBlock(16, 2, feedforward_dim=64, attn_type="efficientkan")
    save_case("block_efficientkan", blk, xin((3, 5, 16), 31))

    torch.manual_seed(32)
    blk = ref_model.TransformerBlock(16, 2, feedforward_dim=64, attn_type="efficientkan")
    blk.ff = torch.nn.Sequential(KANLinear(16, 64), KANLinear(64, 16))  # KAN-MLP extension
    save_case("block_efficientkan_kanmlp", blk, xin((3, 5, 16), 33))

    # ---------------- ViT ----------------
    class ChebyShim(torch.nn.Module):  # labelled shim, harness only
        def __init__(self, inner):
            super().__init__()
            self.inner = inner

        def forward(self, patches):
            return self.inner(patches).view(patches.shape[0], patches.shape[1], -1)

    for kind in ("vanilla", "efficientkan", "fast", "sine", "cheby"):
        torch.manual_seed(40)
        if kind == "cheby":
            vit = ref_model.VisionTransformer((1, 8, 8), n_patches=2, n_blocks=2, d_hidden=16,
                                              n_heads=2, out_d=10, type=kind)
            inner = vit.linear_mapper
            vit.linear_mapper = ChebyShim(inner)
            sd_fix = lambda k: k.replace("linear_mapper.inner.", "linear_mapper.")  # noqa
        else:
            vit = ref_model.VisionTransformer((1, 8, 8), n_patches=2, n_blocks=2, d_hidden=16,
                                              n_heads=2, out_d=10, type=kind)
            sd_fix = lambda k: k  # noqa
        x = xin((3, 1, 8, 8), 41)
        x = x.clone().requires_grad_(True)
        y = vit(x)
        dy = torch.randn(y.shape, generator=torch.Generator().manual_seed(1234))
        y.backward(dy)
        rec = {"x": _np(x), "dy": _np(dy), "y": _np(y), "gx": _np(x.grad),
               "pos_embeddings": _np(vit.pos_embeddings)}
        for k, v in vit.state_dict().items():
            rec["sd/" + sd_fix(k)] = _np(v)
        for k, p in vit.named_parameters():
            if p.grad is not None:
                rec["g/" + sd_fix(k)] = _np(p.grad)
        np.savez_compressed(os.path.join(OUT, f"vit_{kind}.npz"), **rec)
        print(f"vit_{kind}: y{tuple(y.shape)}")

    # ---------------- bit-exact layout pieces ----------------
    vit = ref_model.VisionTransformer((3, 32, 32), n_patches=4, n_blocks=1, d_hidden=24, n_heads=2)
    img = xin((2, 3, 32, 32), 50)
    vit7 = ref_model.VisionTransformer((1, 28, 28), n_patches=7, n_blocks=1, d_hidden=64, n_heads=2)
    img7 = xin((2, 1, 28, 28), 51)
    np.savez_compressed(
        os.path.join(OUT, "layout.npz"),
        img_3x32=_np(img), patches_3x32_np4=_np(vit.patchify(img, 4)),
        img_1x28=_np(img7), patches_1x28_np7=_np(vit7.patchify(img7, 7)),
        pos_17x24=_np(vit.pos_embeddings), pos_50x64=_np(vit7.pos_embeddings),
    )
    # state_dict key/shape lists for the BASELINE configs (host-logic tests)
    import json
    keys = {}
    for name, args in {
        "cfg1_efficientkan": dict(chw=(1, 28, 28), n_patches=7, n_blocks=4, d_hidden=64, n_heads=2, out_d=10, type="efficientkan"),
        "cfg3_fast": dict(chw=(3, 32, 32), n_patches=4, n_blocks=6, d_hidden=192, n_heads=8, out_d=10, type="fast"),
        "cfg3_sine": dict(chw=(3, 32, 32), n_patches=4, n_blocks=6, d_hidden=192, n_heads=8, out_d=10, type="sine"),
        "cfg2_cheby": dict(chw=(1, 28, 28), n_patches=7, n_blocks=4, d_hidden=64, n_heads=2, out_d=10, type="cheby"),
        "vanilla": dict(chw=(1, 28, 28), n_patches=7, n_blocks=2, d_hidden=64, n_heads=2, out_d=10, type="vanilla"),
    }.items():
        torch.manual_seed(0)
        v = ref_model.VisionTransformer(**args)
        keys[name] = {k: list(t.shape) for k, t in v.state_dict().items()}
    with open(os.path.join(OUT, "state_dict_keys.json"), "w") as f:
        json.dump(keys, f, indent=0, sort_keys=True)
    flash_cases()
    maintenance_case()
    print("done")


if __name__ == "__main__":
    main()
