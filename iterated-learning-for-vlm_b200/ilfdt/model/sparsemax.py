"""Drop-in for prototype/model/sparsemax.py: Sparsemax(dim=None).forward(x) projects along `dim`."""
import torch.nn as nn

from ..functional import SparsemaxFunction


class Sparsemax(nn.Module):
    def __init__(self, dim=None):
        super().__init__()
        self.dim = -1 if dim is None else dim   # sparsemax.py:20

    def forward(self, input):
        x = input.movedim(self.dim, -1)
        lead = x.shape
        out = SparsemaxFunction.apply(x.reshape(-1, lead[-1])).reshape(lead).movedim(-1, self.dim)
        self.output = out                        # the reference keeps it too (sparsemax.py:63)
        return out
