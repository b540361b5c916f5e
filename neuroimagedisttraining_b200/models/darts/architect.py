import torch


class Architect:
    """Updates the architecture parameters on validation data: first order (plain gradient) or the second-order
    approximation of DARTS (virtual weight step + finite-difference Hessian-vector product)."""

    def __init__(self, model, arch_lr=3e-4, arch_wd=1e-3, momentum=0.9, weight_decay=3e-4):
        self.model, self.momentum, self.wd = model, momentum, weight_decay
        self.optimizer = torch.optim.Adam(model.arch_parameters(), lr=arch_lr, betas=(0.5, 0.999), weight_decay=arch_wd)

    def step(self, x_train, y_train, x_valid, y_valid, eta, unrolled=False):
        self.optimizer.zero_grad()
        if not unrolled:
            loss = self.model._loss(x_valid, y_valid)
            grads = torch.autograd.grad(loss, self.model.arch_parameters())
        else:
            grads = self._unrolled(x_train, y_train, x_valid, y_valid, eta)
        for p, g in zip(self.model.arch_parameters(), grads):
            p.grad = g.detach().clone()
        self.optimizer.step()

    def _unrolled(self, x_train, y_train, x_valid, y_valid, eta):
        w = self.model.weight_parameters()
        loss = self.model._loss(x_train, y_train)
        gw = torch.autograd.grad(loss, w)
        backup = [p.data.clone() for p in w]
        with torch.no_grad():
            for p, g in zip(w, gw):
                p.sub_(eta * (g + self.wd * p))                      # w' = w - eta * dL_train/dw
        vloss = self.model._loss(x_valid, y_valid)
        d = torch.autograd.grad(vloss, self.model.arch_parameters() + w)
        dalpha, dw = d[:2], d[2:]
        r = 1e-2 / (torch.cat([v.flatten() for v in dw]).norm() + 1e-12)

        def arch_grad_at(sign):
            with torch.no_grad():
                for p, b, v in zip(w, backup, dw):
                    p.copy_(b + sign * r * v)
            return torch.autograd.grad(self.model._loss(x_train, y_train), self.model.arch_parameters())
        gp, gn = arch_grad_at(+1), arch_grad_at(-1)
        with torch.no_grad():
            for p, b in zip(w, backup):
                p.copy_(b)
        return [da - eta * (p - n) / (2 * r) for da, p, n in zip(dalpha, gp, gn)]
