"""GRU module with nn.GRU's parameter names, shapes, init order and (output, h_n) contract, running on
libtt_b200's kernels (reference call sites: simple_two_tower.py:17-32, margin_two_tower.py:12-28,
enhanced_two_tower.py:17-33).  batch_first only; zero initial state."""
import math

import torch
import torch.nn as nn

from . import functional as Fn


class GRU(nn.Module):
    def __init__(self, input_size, hidden_size, num_layers=1, batch_first=True, bidirectional=False, dropout=0.0,
                 bias=True):
        super().__init__()
        if not batch_first or not bias:
            raise NotImplementedError("two_towers_b200.GRU supports batch_first=True, bias=True (all the reference uses)")
        self.input_size, self.hidden_size, self.num_layers = input_size, hidden_size, num_layers
        self.bidirectional, self.dropout, self.batch_first = bidirectional, float(dropout), True
        self._names = []
        for layer in range(num_layers):
            in_l = input_size if layer == 0 else hidden_size * (2 if bidirectional else 1)
            for sfx in ("", "_reverse")[: 2 if bidirectional else 1]:
                for nme, shape in (("weight_ih", (3 * hidden_size, in_l)), ("weight_hh", (3 * hidden_size, hidden_size)),
                                   ("bias_ih", (3 * hidden_size,)), ("bias_hh", (3 * hidden_size,))):
                    full = f"{nme}_l{layer}{sfx}"
                    setattr(self, full, nn.Parameter(torch.empty(*shape)))
                    self._names.append(full)
        self.reset_parameters()
        self._table, self._pad_id = None, -1

    def reset_parameters(self):
        # same draw order as nn.GRU.reset_parameters, so torch.manual_seed(s) reproduces the reference init
        stdv = 1.0 / math.sqrt(self.hidden_size) if self.hidden_size > 0 else 0
        for p in self.parameters():
            nn.init.uniform_(p, -stdv, stdv)

    def flat_weights(self):
        return [getattr(self, n) for n in self._names]

    def forward(self, x, lengths=None, table=None, pad_id=-1, need_output=True):
        """x: float [B,T,input_size] or integer ids [B,T] (needs `table` [V,input_size]).
        Returns (output [B,T,ndir*H], h_n [num_layers*ndir,B,H]) like nn.GRU(batch_first=True); with
        need_output=False (what the towers use: they only read h_n) output is an empty tensor and is never written."""
        return Fn.gru_stack(x, self.flat_weights(), self.input_size, self.hidden_size, self.num_layers,
                            self.bidirectional, self.dropout, self.training, lengths=lengths, table=table,
                            pad_id=pad_id, need_out=need_output)

    def extra_repr(self):
        return (f"{self.input_size}, {self.hidden_size}, num_layers={self.num_layers}, batch_first=True, "
                f"bidirectional={self.bidirectional}, dropout={self.dropout}")
