import torch
import torch.nn as nn
import torch.nn.functional as F

from .genotypes import Genotype, PRIMITIVES
from .operations import OPS, FactorizedReduce, ReLUConvBN


class MixedOp(nn.Module):
    def __init__(self, C, stride):
        super().__init__()
        self._ops = nn.ModuleList()
        for prim in PRIMITIVES:
            op = OPS[prim](C, stride, False)
            if "pool" in prim:
                op = nn.Sequential(op, nn.BatchNorm2d(C, affine=False))
            self._ops.append(op)

    def forward(self, x, weights):
        return sum(w * op(x) for w, op in zip(weights, self._ops))


class Cell(nn.Module):
    def __init__(self, steps, multiplier, C_pp, C_p, C, reduction, reduction_prev):
        super().__init__()
        self.reduction, self._steps, self._multiplier = reduction, steps, multiplier
        self.preprocess0 = FactorizedReduce(C_pp, C, affine=False) if reduction_prev else ReLUConvBN(C_pp, C, 1, 1, 0, affine=False)
        self.preprocess1 = ReLUConvBN(C_p, C, 1, 1, 0, affine=False)
        self._ops = nn.ModuleList()
        for i in range(steps):
            for j in range(2 + i):
                self._ops.append(MixedOp(C, 2 if reduction and j < 2 else 1))

    def forward(self, s0, s1, weights):
        states = [self.preprocess0(s0), self.preprocess1(s1)]
        offset = 0
        for _ in range(self._steps):
            s = sum(self._ops[offset + j](h, weights[offset + j]) for j, h in enumerate(states))
            offset += len(states)
            states.append(s)
        return torch.cat(states[-self._multiplier:], dim=1)


class Network(nn.Module):
    """Searchable network: every edge is a softmax mixture over PRIMITIVES; ``arch_parameters()`` are the alphas."""

    def __init__(self, C=16, num_classes=10, layers=8, criterion=None, steps=4, multiplier=4, stem_multiplier=3):
        super().__init__()
        self._C, self._num_classes, self._layers, self._steps, self._multiplier = C, num_classes, layers, steps, multiplier
        self._criterion = criterion or nn.CrossEntropyLoss()
        C_curr = stem_multiplier * C
        self.stem = nn.Sequential(nn.Conv2d(3, C_curr, 3, padding=1, bias=False), nn.BatchNorm2d(C_curr))
        C_pp, C_p, C_curr = C_curr, C_curr, C
        self.cells = nn.ModuleList()
        reduction_prev = False
        for i in range(layers):
            reduction = i in [layers // 3, 2 * layers // 3]
            if reduction:
                C_curr *= 2
            cell = Cell(steps, multiplier, C_pp, C_p, C_curr, reduction, reduction_prev)
            reduction_prev = reduction
            self.cells.append(cell)
            C_pp, C_p = C_p, multiplier * C_curr
        self.global_pooling = nn.AdaptiveAvgPool2d(1)
        self.classifier = nn.Linear(C_p, num_classes)
        k = sum(2 + i for i in range(steps))
        self.alphas_normal = nn.Parameter(1e-3 * torch.randn(k, len(PRIMITIVES)))
        self.alphas_reduce = nn.Parameter(1e-3 * torch.randn(k, len(PRIMITIVES)))

    def arch_parameters(self):
        return [self.alphas_normal, self.alphas_reduce]

    def weight_parameters(self):
        ids = {id(p) for p in self.arch_parameters()}
        return [p for p in self.parameters() if id(p) not in ids]

    def new(self):
        m = Network(self._C, self._num_classes, self._layers, self._criterion, self._steps, self._multiplier).to(self.alphas_normal.device)
        for x, y in zip(m.arch_parameters(), self.arch_parameters()):
            x.data.copy_(y.data)
        return m

    def forward(self, x):
        s0 = s1 = self.stem(x)
        for cell in self.cells:
            w = F.softmax(self.alphas_reduce if cell.reduction else self.alphas_normal, dim=-1)
            s0, s1 = s1, cell(s0, s1, w)
        return self.classifier(self.global_pooling(s1).flatten(1))

    def _loss(self, x, target):
        return self._criterion(self(x), target)

    def genotype(self) -> Genotype:
        def parse(weights):
            gene, n, start = [], 2, 0
            none = PRIMITIVES.index("none")
            for i in range(self._steps):
                W = weights[start:start + n]
                edges = sorted(range(i + 2), key=lambda e: -max(W[e][k] for k in range(len(W[e])) if k != none))[:2]
                for j in edges:
                    k_best = max((k for k in range(len(W[j])) if k != none), key=lambda k: W[j][k])
                    gene.append((PRIMITIVES[k_best], j))
                start += n
                n += 1
            return gene
        concat = list(range(2 + self._steps - self._multiplier, self._steps + 2))
        return Genotype(normal=parse(F.softmax(self.alphas_normal, -1).tolist()), normal_concat=concat,
                        reduce=parse(F.softmax(self.alphas_reduce, -1).tolist()), reduce_concat=concat)
