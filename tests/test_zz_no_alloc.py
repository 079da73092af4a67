"""kv_kan_fwd_noalloc (include/kanvit_b200.h): pack + exact fp32 forward of one KAN layer with the packed weights in the
library's static device scratch -- no cudaMalloc anywhere on the call.  It has no counterpart in the reference; the
check is against the existing entry points (kv_kan_pack + kv_kan_fwd with caller-allocated buffers), called through the
C ABI, on calls in a row that share the scratch.  Tolerance: the fp32 path's 1e-5 (the two routes run the same kernels,
so the results are in fact equal).  Collected last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _layers(dev):
    import kan_vit_b200 as K
    torch.manual_seed(7)
    return [K.KANLinear(48, 80).to(dev),                      # 9 bases, knots as prologue parameters
            K.ChebyKANLayer(32, 16, 4).to(dev),               # smaller than the call before it: stale scratch behind it
            K.SineKANLayer(24, 40, grid_size=6).to(dev),      # prologue parameters and a bias
            torch.nn.Linear(64, 10).to(dev)]


def _both_routes(m, T, dev):
    """(y from kv_kan_pack + kv_kan_fwd, y from kv_kan_fwd_noalloc, free-memory delta around the second route)."""
    from kan_vit_b200 import _lib as L
    from kan_vit_b200._dispatch import make_spec, slots_of
    lib = L.load()
    spec = make_spec([m])
    flat = slots_of(m)
    x = torch.randn(T, spec.in_features, device=dev)
    flags = spec.dyn_flags() if spec.dyn_flags is not None else 0
    d = spec.desc(T, spec.in_features, spec.out_features, extra_flags=flags)      # no KV_FLAG_BF16_TC: exact kernels
    st = torch.cuda.current_stream().cuda_stream
    pk = spec.packed(flat, dev, None, flags)
    y_ref = torch.empty(T, spec.out_features, device=dev)
    L.check(lib.kv_kan_fwd(C.byref(d), x.data_ptr(), L.ptr(pk["wp"]), None, L.ptr(pk["fp"]), L.ptr(pk["bias"]), None,
                           None, None, None, y_ref.data_ptr(), st), "kv_kan_fwd")
    y = torch.full((T, spec.out_features), float("nan"), device=dev)
    tbl = pk["tbl"]
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    L.check(lib.kv_kan_fwd_noalloc(C.byref(d), tbl.data_ptr(), x.data_ptr(), None, None, None, None, y.data_ptr(), st),
            "kv_kan_fwd_noalloc")
    torch.cuda.synchronize()
    return y_ref, y, free0 - torch.cuda.mem_get_info()[0]


def test_noalloc_forward_matches_pack_plus_forward_on_calls_in_a_row():
    from conftest import relerr
    dev = torch.device("cuda:0")
    layers = _layers(dev)
    # two rounds over the layers: every call inherits the scratch of a different layer; the second round also runs with
    # every kernel already loaded, so the device's free memory must not move at all
    for rnd in range(2):
        for m in layers:
            y_ref, y, used = _both_routes(m, 300 + 17 * rnd, dev)
            assert torch.isfinite(y).all()
            assert relerr(y, y_ref) < TOL, (type(m).__name__, rnd)
            if rnd == 1:
                assert used == 0, (type(m).__name__, used)


def test_noalloc_rejects_what_does_not_fit():
    from kan_vit_b200 import _lib as L
    from kan_vit_b200.ops import KanSpec
    lib = L.load()
    dev = torch.device("cuda:0")
    dummy = torch.zeros(64, device=dev)
    big = KanSpec(L.KV_EFFKAN, 768, 3072, grid_size=5, spline_order=3)       # 21 M packed weights > 1 Mi floats
    d = big.desc(8, 768, 3072)
    rc = lib.kv_kan_fwd_noalloc(C.byref(d), dummy.data_ptr(), dummy.data_ptr(), None, None, None, None,
                                dummy.data_ptr(), None)
    assert rc == -2                                                          # KV_ERR_BAD_SHAPE, nothing launched
    msg = lib.kv_last_cuda_error().decode()
    assert "scratch floats" in msg and "1048576" in msg
    small = KanSpec(L.KV_CHEBY, 16, 16, grid_size=3)
    d = small.desc(8, 16, 16, extra_flags=L.KV_FLAG_BF16_TC)
    assert lib.kv_kan_fwd_noalloc(C.byref(d), dummy.data_ptr(), dummy.data_ptr(), None, None, None, None,
                                  dummy.data_ptr(), None) == -5              # KV_ERR_UNSUPPORTED
    torch.cuda.synchronize()
