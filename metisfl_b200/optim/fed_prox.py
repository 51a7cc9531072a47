"""FedProx (perturbed gradient descent) as a ``torch.optim.Optimizer`` for user-owned training loops.

``w <- w - lr * (g + mu * (w - w*))`` where ``w*`` ("vstar") is the community model the task started from
(reference ``models/keras/optimizers/fed_prox.py:49-58``).  The framework-owned loop does not use this class: there
the same rule is one branch of the fused flat-arena optimizer kernel (K2, ``csrc/kernels/optim.cu``), selected by
``OptimizerConfig.fed_prox``.  ``set_anchor()`` re-captures ``w*`` (call it after installing a new community model).
"""
import torch


class FedProx(torch.optim.Optimizer):
    def __init__(self, params, learning_rate=0.01, proximal_term=0.001):
        if learning_rate < 0 or proximal_term < 0:
            raise ValueError("learning_rate and proximal_term must be non-negative")
        super().__init__(params, dict(lr=learning_rate, mu=proximal_term))
        self.set_anchor()

    @torch.no_grad()
    def set_anchor(self):
        for group in self.param_groups:
            for p in group["params"]:
                self.state[p]["vstar"] = p.detach().clone()

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            ps = [p for p in group["params"] if p.grad is not None]
            if not ps:
                continue
            gs = [p.grad for p in ps]
            vs = [self.state[p]["vstar"] for p in ps]
            # update = lr * g + lr * mu * (w - w*)
            diff = torch._foreach_sub(ps, vs)
            torch._foreach_mul_(diff, group["lr"] * group["mu"])
            torch._foreach_add_(diff, gs, alpha=group["lr"])
            torch._foreach_sub_(ps, diff)
        return loss
