import torch
import torch.nn as nn

from .operations import OPS, FactorizedReduce, Identity, ReLUConvBN


class Cell(nn.Module):
    def __init__(self, genotype, C_pp, C_p, C, reduction, reduction_prev):
        super().__init__()
        self.preprocess0 = FactorizedReduce(C_pp, C) if reduction_prev else ReLUConvBN(C_pp, C, 1, 1, 0)
        self.preprocess1 = ReLUConvBN(C_p, C, 1, 1, 0)
        names, idx = zip(*(genotype.reduce if reduction else genotype.normal))
        self._concat = list(genotype.reduce_concat if reduction else genotype.normal_concat)
        self._steps, self.multiplier, self._indices = len(names) // 2, len(self._concat), idx
        self._ops = nn.ModuleList(OPS[n](C, 2 if reduction and i < 2 else 1, True) for n, i in zip(names, idx))

    def forward(self, s0, s1, drop_prob=0.0):
        states = [self.preprocess0(s0), self.preprocess1(s1)]
        for i in range(self._steps):
            hs = []
            for k in (2 * i, 2 * i + 1):
                h = self._ops[k](states[self._indices[k]])
                if self.training and drop_prob > 0 and not isinstance(self._ops[k], Identity):
                    mask = torch.bernoulli(torch.full((h.size(0), 1, 1, 1), 1 - drop_prob, device=h.device))
                    h = h / (1 - drop_prob) * mask
                hs.append(h)
            states.append(hs[0] + hs[1])
        return torch.cat([states[i] for i in self._concat], dim=1)


class NetworkCIFAR(nn.Module):
    """Discrete network built from a :class:`Genotype` (optionally with the auxiliary head at 2/3 depth)."""

    def __init__(self, C=36, num_classes=10, layers=20, auxiliary=False, genotype=None, drop_path_prob=0.0):
        super().__init__()
        from .genotypes import DARTS_V2
        genotype = genotype or DARTS_V2
        self._layers, self._auxiliary, self.drop_path_prob = layers, auxiliary, drop_path_prob
        C_curr = 3 * C
        self.stem = nn.Sequential(nn.Conv2d(3, C_curr, 3, padding=1, bias=False), nn.BatchNorm2d(C_curr))
        C_pp, C_p, C_curr = C_curr, C_curr, C
        self.cells = nn.ModuleList()
        reduction_prev = False
        for i in range(layers):
            reduction = i in [layers // 3, 2 * layers // 3]
            if reduction:
                C_curr *= 2
            cell = Cell(genotype, C_pp, C_p, C_curr, reduction, reduction_prev)
            reduction_prev = reduction
            self.cells.append(cell)
            C_pp, C_p = C_p, cell.multiplier * C_curr
            if i == 2 * layers // 3:
                C_aux = C_p
        if auxiliary:
            self.auxiliary_head = nn.Sequential(nn.ReLU(inplace=True), nn.AvgPool2d(5, 3, 0, count_include_pad=False), nn.Conv2d(C_aux, 128, 1, bias=False),
                                                nn.BatchNorm2d(128), nn.ReLU(inplace=True), nn.Conv2d(128, 768, 2, bias=False), nn.BatchNorm2d(768),
                                                nn.ReLU(inplace=True), nn.Flatten(), nn.Linear(768, num_classes))
        self.global_pooling = nn.AdaptiveAvgPool2d(1)
        self.classifier = nn.Linear(C_p, num_classes)

    def forward(self, x):
        aux = None
        s0 = s1 = self.stem(x)
        for i, cell in enumerate(self.cells):
            s0, s1 = s1, cell(s0, s1, self.drop_path_prob)
            if i == 2 * self._layers // 3 and self._auxiliary and self.training:
                aux = self.auxiliary_head(s1)
        out = self.classifier(self.global_pooling(s1).flatten(1))
        return (out, aux) if (self._auxiliary and self.training) else out


def _aux_head(C, num_classes, pool_stride, first_pad_free=True):
    """Auxiliary classifier at 2/3 depth: 8x8 inputs for CIFAR (pool stride 3), 14x14 for ImageNet (stride 2)."""
    return nn.Sequential(nn.ReLU(inplace=True), nn.AvgPool2d(5, pool_stride, 0, count_include_pad=False), nn.Conv2d(C, 128, 1, bias=False),
                         nn.BatchNorm2d(128), nn.ReLU(inplace=True), nn.Conv2d(128, 768, 2, bias=False), nn.BatchNorm2d(768),
                         nn.ReLU(inplace=True), nn.Flatten(), nn.Linear(768, num_classes))


class NetworkImageNet(nn.Module):
    """ImageNet-resolution variant (``darts/model.py:161-216``): two stride-2 stems bring 224x224 down to 28x28 before the cells,
    the first cell therefore sees a reduced ``s0``; 7x7 average pool before the classifier."""

    def __init__(self, C=48, num_classes=1000, layers=14, auxiliary=False, genotype=None, drop_path_prob=0.0):
        super().__init__()
        from .genotypes import DARTS_V2
        genotype = genotype or DARTS_V2
        self._layers, self._auxiliary, self.drop_path_prob = layers, auxiliary, drop_path_prob
        self.stem0 = nn.Sequential(nn.Conv2d(3, C // 2, 3, 2, 1, bias=False), nn.BatchNorm2d(C // 2), nn.ReLU(inplace=True),
                                   nn.Conv2d(C // 2, C, 3, 2, 1, bias=False), nn.BatchNorm2d(C))
        self.stem1 = nn.Sequential(nn.ReLU(inplace=True), nn.Conv2d(C, C, 3, 2, 1, bias=False), nn.BatchNorm2d(C))
        C_pp, C_p, C_curr = C, C, C
        self.cells = nn.ModuleList()
        reduction_prev = True
        for i in range(layers):
            reduction = i in [layers // 3, 2 * layers // 3]
            if reduction:
                C_curr *= 2
            cell = Cell(genotype, C_pp, C_p, C_curr, reduction, reduction_prev)
            reduction_prev = reduction
            self.cells.append(cell)
            C_pp, C_p = C_p, cell.multiplier * C_curr
            if i == 2 * layers // 3:
                C_aux = C_p
        if auxiliary:
            self.auxiliary_head = _aux_head(C_aux, num_classes, 2)
        self.global_pooling = nn.AdaptiveAvgPool2d(1)        # == AvgPool2d(7) at 224x224, and still defined at other resolutions
        self.classifier = nn.Linear(C_p, num_classes)

    def forward(self, x):
        aux = None
        s0 = self.stem0(x)
        s1 = self.stem1(s0)
        for i, cell in enumerate(self.cells):
            s0, s1 = s1, cell(s0, s1, self.drop_path_prob)
            if i == 2 * self._layers // 3 and self._auxiliary and self.training:
                aux = self.auxiliary_head(s1)
        out = self.classifier(self.global_pooling(s1).flatten(1))
        return (out, aux) if (self._auxiliary and self.training) else out
