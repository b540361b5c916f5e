"""GDAS-style search network: every forward pass samples ONE operation per edge with a straight-through Gumbel-softmax
(temperature ``tau``, annealed by the search loop), so only the sampled branch is evaluated.  Reference:
``fedml_api/model/cv/darts/model_search_gdas.py:9-188`` (``Network_GumbelSoftmax``)."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .genotypes import PRIMITIVES
from .model_search import Cell as _SoftCell, MixedOp as _SoftMixedOp, Network as _SoftNetwork

CONV_PRIMITIVES = tuple(i for i, p in enumerate(PRIMITIVES) if "conv" in p)


class MixedOp(_SoftMixedOp):
    """Evaluates only the branches whose (one-hot) weight is non-zero; the weight stays in the graph so the straight-through
    estimator still reaches the architecture parameters."""

    def forward(self, x, weights, active=None):
        idx = active if active is not None else [j for j, w in enumerate(weights.tolist()) if abs(w) > 1e-10]
        outs = [weights[j] * self._ops[j](x) for j in idx]
        return outs[0] if len(outs) == 1 else sum(outs)


class Cell(_SoftCell):
    MIXED = MixedOp

    def __init__(self, steps, multiplier, C_pp, C_p, C, reduction, reduction_prev):
        super().__init__(steps, multiplier, C_pp, C_p, C, reduction, reduction_prev)
        self._ops = nn.ModuleList(MixedOp(C, 2 if reduction and j < 2 else 1) for i in range(steps) for j in range(2 + i))

    def forward(self, s0, s1, weights):
        states = [self.preprocess0(s0), self.preprocess1(s1)]
        active = [[j for j, w in enumerate(row) if abs(w) > 1e-10] for row in weights.tolist()]    # one host read per cell
        offset = 0
        for _ in range(self._steps):
            s = sum(self._ops[offset + j](h, weights[offset + j], active[offset + j]) for j, h in enumerate(states))
            offset += len(states)
            states.append(s)
        return torch.cat(states[-self._multiplier:], dim=1)


class NetworkGumbelSoftmax(_SoftNetwork):
    """Same skeleton and alphas as the softmax-relaxed :class:`~.model_search.Network`; the edge weights are hard one-hot samples."""

    def __init__(self, C=16, num_classes=10, layers=8, criterion=None, steps=4, multiplier=4, stem_multiplier=3, tau=5.0):
        super().__init__(C, num_classes, layers, criterion, steps, multiplier, stem_multiplier)
        cells, C_pp, C_p, C_curr, reduction_prev = [], stem_multiplier * C, stem_multiplier * C, C, False
        for i in range(layers):
            reduction = i in [layers // 3, 2 * layers // 3]
            if reduction:
                C_curr *= 2
            cells.append(Cell(steps, multiplier, C_pp, C_p, C_curr, reduction, reduction_prev))
            reduction_prev = reduction
            C_pp, C_p = C_p, multiplier * C_curr
        self.cells = nn.ModuleList(cells)
        self.tau = tau

    def set_tau(self, tau):
        self.tau = tau

    def get_tau(self):
        return self.tau

    def new(self):
        m = NetworkGumbelSoftmax(self._C, self._num_classes, self._layers, self._criterion, self._steps, self._multiplier, tau=self.tau)
        m = m.to(self.alphas_normal.device)
        for x, y in zip(m.arch_parameters(), self.arch_parameters()):
            x.data.copy_(y.data)
        return m

    def forward(self, x):
        s0 = s1 = self.stem(x)
        for cell in self.cells:
            alphas = self.alphas_reduce if cell.reduction else self.alphas_normal
            s0, s1 = s1, cell(s0, s1, F.gumbel_softmax(alphas, self.tau, hard=True))
        return self.classifier(self.global_pooling(s1).flatten(1))

    def genotype_with_counts(self):
        """(genotype, number of convolutional ops chosen in the normal cell, ... in the reduction cell) -- the reference's
        ``genotype()`` return value for this class (``model_search_gdas.py:142-188``)."""
        g = self.genotype()
        conv = {PRIMITIVES[i] for i in CONV_PRIMITIVES}
        return g, sum(op in conv for op, _ in g.normal), sum(op in conv for op, _ in g.reduce)


Network_GumbelSoftmax = NetworkGumbelSoftmax
