"""TransformerBlock / VisionTransformer -- drop-in for model.py:14-169 of the reference."""
import numpy
import torch
import torch.nn as nn

from . import ops
from ._dispatch import make_spec, slots_of
from .attention import MSA, FlashAttention
from .models.cheby import ChebyKANLayer
from .models.effkan import KANLinear
from .models.fastkan import FastKANLayer
from .models.nfkan import NaiveFourierKANLayer
from .models.sinekan import SineKANLayer


class TransformerBlock(nn.Module):
    """x += MSA(LN(x)); x += ff(LN(x)).  ``kan_mlp=True`` (keyword-only extension, SURVEY.md
    section 0.2) builds ``ff`` from two KANLinear layers instead of Linear/ReLU/Linear."""

    def __init__(self, d_model, n_heads, feedforward_dim=128, attn_type="vanilla", *, kan_mlp=False):
        super().__init__()
        self.norm1 = nn.LayerNorm(d_model)
        self.attn = MSA(d_model, n_heads, type=attn_type)
        self.norm2 = nn.LayerNorm(d_model)
        self.kan_mlp = kan_mlp
        if kan_mlp:
            self.ff = nn.Sequential(KANLinear(d_model, feedforward_dim), KANLinear(feedforward_dim, d_model))
        else:
            self.ff = nn.Sequential(nn.Linear(d_model, feedforward_dim), nn.ReLU(inplace=True),
                                    nn.Linear(feedforward_dim, d_model))
        self._specs = None

    def forward2d(self, x, B, N):
        if self._specs is None:
            self._specs = (make_spec([self.ff[0]]), make_spec([self.ff[-1]]))
        h, skip = ops.layer_norm_fork(x, self.norm1.weight, self.norm1.bias, self.norm1.eps)
        x = self.attn.forward2d(h, B, N, residual=skip)
        h, skip = ops.layer_norm_fork(x, self.norm2.weight, self.norm2.bias, self.norm2.eps)
        h = ops.kan_apply(h, self._specs[0], [slots_of(self.ff[0])], relu=not self.kan_mlp)
        return ops.kan_apply(h, self._specs[1], [slots_of(self.ff[-1])], residual=skip)

    def forward(self, x):
        B, N, d = x.shape
        return self.forward2d(x.reshape(B * N, d).contiguous().float(), B, N).reshape(B, N, d)


class VisionTransformer(nn.Module):
    def __init__(self, chw, n_patches=7, n_blocks=4, d_hidden=64, n_heads=2, out_d=10,
                 type: str = "vanilla", *, kan_mlp=False):
        super().__init__()
        self.chw = chw
        self.n_patches = n_patches
        self.n_blocks = n_blocks
        self.n_heads = n_heads
        self.d_hidden = d_hidden
        assert chw[1] % n_patches == 0
        assert chw[2] % n_patches == 0
        self.patch_size = (chw[1] // n_patches, chw[2] // n_patches)
        self.input_d = int(chw[0] * self.patch_size[0] * self.patch_size[1])

        if type == "vanilla" or type == "flash-attn":
            self.linear_mapper = nn.Linear(self.input_d, d_hidden)
        elif type == "efficientkan":
            self.linear_mapper = KANLinear(self.input_d, d_hidden)
        elif type == "sine":
            self.linear_mapper = SineKANLayer(self.input_d, d_hidden, grid_size=28)
        elif type == "fourier":
            # the reference passes a misspelt kwarg here and cannot construct this variant
            # (model.py:74 vs models/nfkan.py:6); the intended layer is built instead
            self.linear_mapper = NaiveFourierKANLayer(self.input_d, d_hidden, 28)
        elif type == "cheby":
            self.linear_mapper = ChebyKANLayer(self.input_d, d_hidden, 4)
        elif type == "fast":
            self.linear_mapper = FastKANLayer(self.input_d, d_hidden)
        else:
            raise ValueError(f"Unknown transformer type: {type}")

        self.v_class = nn.Parameter(torch.randn(1, d_hidden))
        self.register_buffer('pos_embeddings', self.positional_embeddings(n_patches ** 2 + 1, d_hidden),
                             persistent=False)
        if type == "flash-attn":
            # bare attention modules: no LayerNorm / residual / MLP around them (model.py:93-95)
            self.blocks = nn.ModuleList([FlashAttention(dim=d_hidden, heads=n_heads) for _ in range(n_blocks)])
        else:
            self.blocks = nn.ModuleList([
                TransformerBlock(d_model=d_hidden, n_heads=n_heads, feedforward_dim=4 * d_hidden,
                                 attn_type=type, kan_mlp=kan_mlp)
                for _ in range(n_blocks)])
        self.mlp_head = nn.Sequential(nn.LayerNorm(d_hidden), nn.Linear(d_hidden, out_d))
        self._specs = None

    def patchify(self, images, n_patches):
        """(B,C,H,W) -> (B, np^2, C*ph*pw) float32; patch index i*np+j, feature c*ph*pw+y*pw+x."""
        n, c, h, w = images.shape
        assert h % n_patches == 0 and w % n_patches == 0
        return ops.patchify(images, n_patches)

    def positional_embeddings(self, seq_length, d):
        """pos[i,j] = sin(i/10000^(j/d)) (j even) | cos(i/10000^(j/d)) (j odd), evaluated in
        float64 and stored float32 (bit-identical to model.py:128-140)."""
        i = numpy.arange(seq_length, dtype=numpy.float64)[:, None]
        j = numpy.arange(d, dtype=numpy.float64)[None, :]
        arg = i / numpy.power(10000.0, j / d)
        even = (numpy.arange(d) % 2 == 0)[None, :]
        return torch.from_numpy(numpy.where(even, numpy.sin(arg), numpy.cos(arg))).to(torch.float32)

    def forward(self, images):
        if self._specs is None:
            self._specs = (make_spec([self.linear_mapper]), make_spec([self.mlp_head[1]]))
        patches = self.patchify(images, self.n_patches)
        B, P2, input_d = patches.shape
        N, d = P2 + 1, self.d_hidden
        x = ops.kan_apply(patches.view(B * P2, input_d), self._specs[0], [slots_of(self.linear_mapper)],
                          pos=self.pos_embeddings, v_class=self.v_class, rows_in=P2, rows_out=N, row_shift=1)
        if isinstance(self.blocks[0], FlashAttention):
            for blk in self.blocks:
                x = blk(x.view(B, N, d)).reshape(B * N, d)
        else:
            for blk in self.blocks:
                x = blk.forward2d(x, B, N)
        cls = x.view(B, N * d)[:, :d]
        head_ln = self.mlp_head[0]
        h = ops.layer_norm(cls, head_ln.weight, head_ln.bias, head_ln.eps)
        return ops.kan_apply(h, self._specs[1], [slots_of(self.mlp_head[1])])
