"""MSA -- drop-in for attention.py:112-202 of the reference.

The reference loops in Python over samples and heads, calling three d_head->d_head maps per
(sample, head).  Every map is row-wise, so here all B*N rows of all heads go through ONE
grouped launch (groups ordered head-major, q/k/v sharing the head's input slice) that writes
[B*N][H][3][d_head], followed by one fused softmax-attention launch.
"""
import torch

from . import _lib as L
from . import ops
from ._dispatch import make_spec, slots_of
from .models.cheby import ChebyKANLayer
from .models.effkan import KANLinear
from .models.fastkan import FastKANLayer
from .models.sinekan import SineKANLayer


class FlashAttention(torch.nn.Module):
    """Drop-in for attention.py:13-109 (the block of ``type="flash-attn"``, model.py:93-95): bias-free
    ``to_q`` / ``to_kv`` / ``to_out`` projections around FlashAttentionFunction (utils.py:137-295).

    The three projections run as Linear GEMM launches and the attention itself (optional ``context``,
    boolean ``mask``, the reference's ``causal`` rule) as one kernel per direction.  The q/k bucket sizes
    only set the summation order of the reference's Python tile loops; they are accepted and ignored.
    ``parallel=True`` cannot run in the reference either (attention.py:91-97 replaces q by a list);
    ``mixed_precision`` is an autocast stub there (attention.py:58-59): use ``set_precision("bf16")``."""

    def __init__(self, *, dim, heads=8, dim_head=64, causal=False, q_bucket_size=512, k_bucket_size=1024,
                 parallel=False, mixed_precision=False):
        super().__init__()
        if parallel:
            raise NotImplementedError("FlashAttention(parallel=True) does not work in the reference "
                                      "(attention.py:91-97); use kan_vit_b200.trainer for data parallelism")
        self.heads = heads
        self.dim_head = dim_head
        self.causal = causal
        self.parallel = parallel
        self.mixed_precision = mixed_precision
        inner_dim = heads * dim_head
        self.to_q = torch.nn.Linear(dim, inner_dim, bias=False)
        self.to_kv = torch.nn.Linear(dim, inner_dim * 2, bias=False)
        self.to_out = torch.nn.Linear(inner_dim, dim, bias=False)
        self.q_bucket_size = q_bucket_size
        self.k_bucket_size = k_bucket_size
        self._specs = None
        self._spec_cat = None

    def forward(self, x, context=None, mask=None, q_bucket_size=None, k_bucket_size=None):
        if self._specs is None:
            self._specs = tuple(make_spec([m]) for m in (self.to_q, self.to_kv, self.to_out))
        B, Nq, dim = x.shape
        ctx = x if context is None else context
        Nk = ctx.shape[1]
        inner = self.heads * self.dim_head
        x2d = x.reshape(B * Nq, dim)
        if (ops.get_precision() == "bf16" and context is None and mask is None and not self.causal
                and ops.attention_tc_ok(Nq, self.heads, self.dim_head)):
            # plain self-attention in bf16 mode: one projection GEMM whose output rows are ordered
            # (head, {q,k,v}, d_head) -- the layout of the tensor-core attention kernels used by MSA
            H, dh = self.heads, self.dim_head
            w = torch.cat([self.to_q.weight.view(H, dh, dim), self.to_kv.weight[:inner].view(H, dh, dim),
                           self.to_kv.weight[inner:].view(H, dh, dim)], dim=1).reshape(3 * inner, dim)
            if self._spec_cat is None:
                self._spec_cat = ops.KanSpec(L.KV_LINEAR, dim, 3 * inner)
            qkv = ops.kan_apply(x2d, self._spec_cat, [[w, None]])
            out = ops.attention(qkv, B, Nq, H, dh)
            return ops.kan_apply(out, self._specs[2], [slots_of(self.to_out)]).reshape(B, Nq, dim)
        q = ops.kan_apply(x2d, self._specs[0], [slots_of(self.to_q)])
        kv = ops.kan_apply(x2d if context is None else ctx.reshape(B * Nk, dim), self._specs[1],
                           [slots_of(self.to_kv)])
        out = ops.attention_general(q, kv[:, :inner], kv[:, inner:], B, Nq, Nk, self.heads, self.dim_head,
                                    mask=mask, causal=self.causal)
        return ops.kan_apply(out, self._specs[2], [slots_of(self.to_out)]).reshape(B, Nq, dim)


class MSA(torch.nn.Module):
    def __init__(self, d, n_heads=4, type: str = "vanilla"):
        super().__init__()
        self.d = d
        self.n_heads = n_heads
        assert d % n_heads == 0
        d_head = int(d / n_heads)

        def maps(make):
            return torch.nn.ModuleList([make() for _ in range(n_heads)])

        makers = {
            "vanilla": lambda: torch.nn.Linear(d_head, d_head),
            "flash-attn": lambda: torch.nn.Linear(d_head, d_head),
            "fourier": lambda: torch.nn.Linear(d_head, d_head),
            "efficientkan": lambda: KANLinear(d_head, d_head),
            "fast": lambda: FastKANLayer(d_head, d_head),
            "sine": lambda: SineKANLayer(d_head, d_head, grid_size=4),
            "cheby": lambda: ChebyKANLayer(d_head, d_head, 4),
        }
        try:
            if type not in makers:
                raise ValueError(f"{type} invalid. Please use a different argument.")
            self.q_mappings = maps(makers[type])
            self.k_mappings = maps(makers[type])
            self.v_mappings = maps(makers[type])
        except Exception:
            # the reference swallows the error here (attention.py:174-176); forward then fails
            print(f"Error initializing MSA with the following arguments: {type}, {d}, {n_heads}.")
        self.d_head = d_head
        self.softmax = torch.nn.Softmax(dim=-1)
        self._spec = None

    def _groups(self):
        mods = []
        for h in range(self.n_heads):
            mods += [self.q_mappings[h], self.k_mappings[h], self.v_mappings[h]]
        return mods

    def forward2d(self, x2d, B, N, residual=None):
        """x2d [B*N, d] -> attention output [B*N, d] (+ residual)."""
        mods = self._groups()
        if self._spec is None:
            self._spec = make_spec(mods, x_share=3)
        slots = [slots_of(m) for m in mods]
        if ops.qkv_attention_fused_ok(self._spec, slots, N, self.n_heads, self.d_head):
            # bf16 tensor-core route: q/k/v GEMM -> bf16 attention image -> tcgen05 attention, one autograd node
            return ops.qkv_attention(x2d, self._spec, slots, B, N, self.n_heads, self.d_head, residual=residual)
        qkv = ops.kan_apply(x2d, self._spec, slots)
        return ops.attention(qkv, B, N, self.n_heads, self.d_head, residual=residual)

    def forward(self, sequences):
        B, N, d = sequences.shape
        assert d == self.d
        return self.forward2d(sequences.reshape(B * N, d), B, N).reshape(B, N, d)
