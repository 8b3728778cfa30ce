"""Per-tensor ``torch.optim.Optimizer`` implementations of the reference update rules.

``Adam`` restates optim.py:162-231 (fairseq Adam: AdamW-style decoupled decay applied *before*
the Adam update, bias correction folded into the step size, eps added outside the sqrt, fp32
state even for fp16 params). ``Adadelta`` restates optim.py:263-304 (coupled L2 decay). They are
the simple path for small CPU models and the in-repo oracle for the flat / fused engines
(hetseq_b200/optim/flat_optimizer.py and csrc/optim_sm100.cu), which must match them.
"""
from __future__ import annotations

import math

import torch


class Adam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad)
        super().__init__(params, defaults)

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            beta1, beta2 = group['betas']
            for p in group['params']:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError('Adam does not support sparse gradients')
                grad = p.grad.float()
                p32 = p.data.float()
                state = self.state[p]
                if len(state) == 0:
                    state['step'] = 0
                    state['exp_avg'] = torch.zeros_like(p32)
                    state['exp_avg_sq'] = torch.zeros_like(p32)
                    if group['amsgrad']:
                        state['max_exp_avg_sq'] = torch.zeros_like(p32)
                else:
                    for k in ('exp_avg', 'exp_avg_sq', 'max_exp_avg_sq'):
                        if k in state:
                            state[k] = state[k].to(p32)
                m, v = state['exp_avg'], state['exp_avg_sq']
                state['step'] += 1
                m.mul_(beta1).add_(grad, alpha=1 - beta1)
                v.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                if group['amsgrad']:
                    torch.maximum(state['max_exp_avg_sq'], v, out=state['max_exp_avg_sq'])
                    denom = state['max_exp_avg_sq'].sqrt().add_(group['eps'])
                else:
                    denom = v.sqrt().add_(group['eps'])
                bc1 = 1 - beta1 ** state['step']
                bc2 = 1 - beta2 ** state['step']
                step_size = group['lr'] * math.sqrt(bc2) / bc1
                if group['weight_decay'] != 0:
                    p32.add_(p32, alpha=-group['weight_decay'] * group['lr'])
                p32.addcdiv_(m, denom, value=-step_size)
                p.data.copy_(p32)
        return loss


class Adadelta(torch.optim.Optimizer):
    def __init__(self, params, lr=1.0, rho=0.9, eps=1e-6, weight_decay=0):
        if not 0.0 <= lr:
            raise ValueError('Invalid learning rate: {}'.format(lr))
        if not 0.0 <= rho <= 1.0:
            raise ValueError('Invalid rho value: {}'.format(rho))
        if not 0.0 <= eps:
            raise ValueError('Invalid epsilon value: {}'.format(eps))
        if not 0.0 <= weight_decay:
            raise ValueError('Invalid weight_decay value: {}'.format(weight_decay))
        super().__init__(params, dict(lr=lr, rho=rho, eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            rho, eps = group['rho'], group['eps']
            for p in group['params']:
                if p.grad is None:
                    continue
                grad = p.grad
                if grad.is_sparse:
                    raise RuntimeError('Adadelta does not support sparse gradients')
                state = self.state[p]
                if len(state) == 0:
                    state['step'] = 0
                    state['square_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state['acc_delta'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                sq, acc = state['square_avg'], state['acc_delta']
                state['step'] += 1
                if group['weight_decay'] != 0:
                    grad = grad.add(p, alpha=group['weight_decay'])
                sq.mul_(rho).addcmul_(grad, grad, value=1 - rho)
                std = sq.add(eps).sqrt_()
                delta = acc.add(eps).sqrt_().div_(std).mul_(grad)
                p.add_(delta, alpha=-group['lr'])
                acc.mul_(rho).addcmul_(delta, delta, value=1 - rho)
        return loss
