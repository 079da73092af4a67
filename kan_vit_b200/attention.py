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
        qkv = ops.kan_apply(x2d, self._spec, [slots_of(m) for m in mods])
        return ops.attention(qkv, B, N, self.n_heads, self.d_head, residual=residual)

    def forward(self, sequences):
        B, N, d = sequences.shape
        assert d == self.d
        return self.forward2d(sequences.reshape(B * N, d), B, N).reshape(B, N, d)
