"""Plain PyTorch references (computed in float64, returned as fp32) of the native ops, on the engine's layouts (NHWC / KRSC).

Used by the numerics tests (every CUDA kernel is compared with the same op here) and as the
documentation of what each kernel computes.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def nhwc_to_nchw(x):
    return x.permute(0, 3, 1, 2)


def nchw_to_nhwc(x):
    return x.permute(0, 2, 3, 1).contiguous()


def krsc_to_oihw(w):
    return w.permute(0, 3, 1, 2)


def conv_fwd(x_nhwc, w_krsc, bias):
    y = F.conv2d(nhwc_to_nchw(x_nhwc.double()), krsc_to_oihw(w_krsc.double()),
                 None if bias is None else bias.double(), stride=1, padding=1)
    return nchw_to_nhwc(y).float()


def conv_dgrad(dy_nhwc, w_krsc):
    dx = F.conv_transpose2d(nhwc_to_nchw(dy_nhwc.double()), krsc_to_oihw(w_krsc.double()), stride=1, padding=1)
    return nchw_to_nhwc(dx).float()


def conv_wgrad(dy_nhwc, x_nhwc):
    x = nhwc_to_nchw(x_nhwc.double())
    dy = nhwc_to_nchw(dy_nhwc.double())
    cout, cin = dy.shape[1], x.shape[1]
    dw = torch.nn.grad.conv2d_weight(x, (cout, cin, 3, 3), dy, stride=1, padding=1)
    return dw.permute(0, 2, 3, 1).contiguous().float()        # OIHW -> KRSC


def channel_stats(y_nhwc):
    y = y_nhwc.float().reshape(-1, y_nhwc.shape[-1])
    return torch.cat([y.sum(0), (y * y).sum(0)])


def bn_relu_pool_fwd(y_nhwc, gamma, beta, running_mean, running_var, pool, training=True, eps=1e-5,
                     momentum=0.1):
    """Returns (out_nhwc fp32, new_running_mean, new_running_var)."""
    x = nhwc_to_nchw(y_nhwc.float())
    rm, rv = running_mean.clone().float(), running_var.clone().float()
    z = F.batch_norm(x, rm, rv, gamma.float(), beta.float(), training, momentum, eps)
    z = F.relu(z)
    if pool:
        z = F.max_pool2d(z, 2, 2)
    return nchw_to_nhwc(z), rm, rv


def bn_relu_pool_bwd(dout_nhwc, y_nhwc, gamma, beta, pool, eps=1e-5):
    """Returns (dy_nhwc, dgamma, dbeta) via autograd on the fp32 forward."""
    y = y_nhwc.float().clone().requires_grad_(True)
    g = gamma.float().clone().requires_grad_(True)
    b = beta.float().clone().requires_grad_(True)
    c = y.shape[-1]
    z = F.batch_norm(nhwc_to_nchw(y), torch.zeros(c, device=y.device), torch.ones(c, device=y.device), g, b, True,
                     0.1, eps)
    z = F.relu(z)
    if pool:
        z = F.max_pool2d(z, 2, 2)
    z.backward(nhwc_to_nchw(dout_nhwc.float()))
    return y.grad, g.grad, b.grad


def fc_ce(act, w, bias, target):
    """Returns (logits, loss, dlogits, dact, dw, db, correct)."""
    a = act.double().clone().requires_grad_(True)
    w_ = w.double().clone().requires_grad_(True)
    b_ = bias.double().clone().requires_grad_(True)
    logits = F.linear(a, w_, b_)
    logits.retain_grad()
    loss = F.cross_entropy(logits, target)
    loss.backward()
    correct = (logits.argmax(1) == target).sum()
    f = lambda t: t.detach().float()
    return f(logits), f(loss), f(logits.grad), f(a.grad), f(w_.grad), f(b_.grad), correct


def sgd(param, mom, grad, lr, momentum, wd, grad_scale=1.0):
    g = grad * grad_scale + wd * param
    mom = momentum * mom + g
    return param - lr * mom, mom
