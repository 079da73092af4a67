"""SineKANLayer -- drop-in for models/sinekan.py:7-109 of the reference."""
import math

import torch

from .. import _lib as L
from ..ops import KanSpec, kan_apply


def forward_step(i_n, grid_size, A, K, C):
    """Phase recurrence of the reference (models/sinekan.py:7-23): (A*g^-K + C) * i_n."""
    return (A * grid_size ** (-K) + C) * i_n


class SineKANLayer(torch.nn.Module):
    """Y[t,o] = sum_i sum_g sin(x[t,i]*freq[g] + phase[i,g]) * amplitudes[o,i,g] + bias[o]."""

    def __init__(self, input_dim, output_dim, device='cpu', grid_size=5, is_first=False, add_bias=True,
                 norm_freq=True):
        super().__init__()
        self.grid_size = grid_size
        self.device = device
        self.is_first = is_first
        self.add_bias = add_bias
        self.input_dim = input_dim
        self.output_dim = output_dim
        self.A, self.K, self.C = 0.9724108095811765, 0.9884401790754128, 0.999449553483052

        self.grid_norm_factor = (torch.arange(grid_size) + 1).reshape(1, 1, grid_size)
        raw = torch.empty(output_dim, input_dim, 1)
        raw = raw.normal_(0, .4) if is_first else raw.uniform_(-1, 1)
        self.amplitudes = torch.nn.Parameter(raw / output_dim / self.grid_norm_factor)

        grid_phase = torch.arange(1, grid_size + 1).reshape(1, 1, 1, grid_size) / (grid_size + 1)
        self.input_phase = torch.linspace(0, math.pi, input_dim).reshape(1, 1, input_dim, 1).to(device)
        phase = grid_phase.to(device) + self.input_phase
        freq = torch.arange(1, grid_size + 1).float().reshape(1, 1, 1, grid_size)
        if norm_freq:
            freq = freq / (grid_size + 1) ** (1 - is_first)
        self.freq = torch.nn.Parameter(freq)
        for i in range(1, self.grid_size):
            phase = forward_step(phase, i, self.A, self.K, self.C)
        self.register_buffer('phase', phase)
        if self.add_bias:
            self.bias = torch.nn.Parameter(torch.ones(1, output_dim) / output_dim)
        self._spec = KanSpec(L.KV_SINE, input_dim, output_dim, grid_size=grid_size)

    def forward_step(self, i_n, grid_size, A, K, C):
        return forward_step(i_n, grid_size, A, K, C)

    def _slots(self):
        return [self.amplitudes, self.freq, self.phase, self.bias if self.add_bias else None]

    def forward(self, x):
        shape = x.shape[:-1] + (self.output_dim,)
        y = kan_apply(x.reshape(-1, self.input_dim), self._spec, [self._slots()])
        return y.reshape(shape)
