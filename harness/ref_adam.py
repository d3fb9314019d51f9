"""Restatement of the reference's sparse Adam (scene/OurAdam.py:105-175,249-337) in the torch ops it
uses -- the reference arm of the benchmark and the checker of tests/ for h3dgs_sparse_adam.
TEST INFRASTRUCTURE: never imported by the product path."""
import math

import torch


class RefSparseAdam:
    def __init__(self, param_groups, betas=(0.9, 0.999), eps=1e-15):
        self.param_groups = [dict(g) for g in param_groups]
        self.betas = betas
        self.eps = eps
        self.state = {}

    def zero_grad(self):
        for g in self.param_groups:
            for p in g["params"]:
                p.grad = None

    @torch.no_grad()
    def step(self, relevant):
        beta1, beta2 = self.betas
        for g in self.param_groups:
            lr = g["lr"]
            for p in g["params"]:
                if p.grad is None:
                    continue
                st = self.state.setdefault(id(p), {})
                if not st:
                    st["step"] = torch.tensor(0.)
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                st["step"] += 1
                step = st["step"].item()
                dense = relevant is None or relevant.size(0) == 0
                if dense:
                    grad, exp_avg, exp_avg_sq, param = p.grad, st["exp_avg"], st["exp_avg_sq"], p
                else:
                    grad = p.grad[relevant]
                    exp_avg = st["exp_avg"][relevant]
                    exp_avg_sq = st["exp_avg_sq"][relevant]
                    param = p[relevant]
                exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                exp_avg_sq.mul_(beta2).addcmul_(grad, grad.conj(), value=1 - beta2)
                bias_correction1 = 1 - beta1 ** step
                bias_correction2 = 1 - beta2 ** step
                step_size = lr / bias_correction1
                bias_correction2_sqrt = math.sqrt(bias_correction2)
                denom = (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(self.eps)
                param.addcdiv_(exp_avg, denom, value=-step_size)
                if not dense:
                    st["exp_avg"][relevant] = exp_avg
                    st["exp_avg_sq"][relevant] = exp_avg_sq
                    p[relevant] = param
