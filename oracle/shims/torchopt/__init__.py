"""Oracle-only restatement of the slice of `torchopt` 0.7.1 the reference calls.

TEST INFRASTRUCTURE.  torchopt==0.7.1 is pinned by the reference
(requirements.txt:132) and is NOT vendored under /root/reference nor installable
here (no network).  Call sites this shim serves: autoencoders/ensemble.py:27
(`torchopt.adam`), :29 (`torchopt.sgd`), :94-95 (`optimizer.init` under vmap),
:123/:182 (`optimizer.update` under vmap), :191 (`torchopt.apply_updates`).

Published algorithm restated (torchopt/transform/scale_by_adam.py,
torchopt/transform/utils.py, torchopt/update.py of the 0.7.1 release), PARITY
UNPINNED against a real wheel - there is none on this machine:

  adam(lr, betas=(.9,.999), eps=1e-8, weight_decay=0, eps_root=0, maximize=False)
    = chain(scale_by_adam(b1,b2,eps,eps_root), scale(-lr))
  init:   mu = zeros_like(p), nu = zeros_like(p), count = int64 zeros(1) per leaf
  update(updates, state, params=None, inplace=True):
      mu.mul_(b1).add_(g, alpha=1-b1)                 # in place  -> persists
      nu.mul_(b2).addcmul_(g, g, value=1-b2)          # in place  -> persists
      count_inc = count + 1                           # OUT of place
      mu_hat = mu.div(1 - b1**count_inc)              # out of place
      nu_hat = nu.div(1 - b2**count_inc)
      upd = mu_hat.div_(nu_hat.add_(eps_root).sqrt_().add(eps))
      upd.mul_(-lr)
      return upd, ScaleByAdamState(mu, nu, count_inc)
  apply_updates(params, updates, inplace=True): p.add_(u)

The in-place / out-of-place split is what makes the reference's step counter
stay at zero (ensemble.py:185-189 copies the new state into a *clone*): bias
correction is evaluated at t=1 on every step.
"""
from typing import Any, Callable, NamedTuple

import torch
import torch.utils._pytree as _pt


class GradientTransformation(NamedTuple):
    init: Callable
    update: Callable


class ScaleByAdamState(NamedTuple):
    mu: Any
    nu: Any
    count: Any


class EmptyState(NamedTuple):
    pass


def _scale_by_adam(b1, b2, eps, eps_root):
    def init_fn(params):
        mu = _pt.tree_map(torch.zeros_like, params)
        nu = _pt.tree_map(torch.zeros_like, params)
        count = _pt.tree_map(lambda t: torch.zeros(1, dtype=torch.int64, device=t.device).squeeze_(), params)
        return ScaleByAdamState(mu=mu, nu=nu, count=count)

    def update_fn(updates, state, *, params=None, inplace=True):
        if inplace:
            mu = _pt.tree_map(lambda g, t: t.mul_(b1).add_(g, alpha=1 - b1), updates, state.mu)
            nu = _pt.tree_map(lambda g, t: t.mul_(b2).addcmul_(g, g, value=1 - b2), updates, state.nu)
        else:
            mu = _pt.tree_map(lambda g, t: t.mul(b1).add(g, alpha=1 - b1), updates, state.mu)
            nu = _pt.tree_map(lambda g, t: t.mul(b2).addcmul(g, g, value=1 - b2), updates, state.nu)
        count_inc = _pt.tree_map(lambda c: c + 1, state.count)
        mu_hat = _pt.tree_map(lambda t, c: t.div(1 - pow(b1, c)), mu, count_inc)
        nu_hat = _pt.tree_map(lambda t, c: t.div(1 - pow(b2, c)), nu, count_inc)
        upd = _pt.tree_map(lambda m, v: m.div_(v.add_(eps_root).sqrt_().add(eps)), mu_hat, nu_hat)
        return upd, ScaleByAdamState(mu=mu, nu=nu, count=count_inc)

    return GradientTransformation(init_fn, update_fn)


def _scale(step_size):
    def init_fn(params):
        return EmptyState()

    def update_fn(updates, state, *, params=None, inplace=True):
        if inplace:
            updates = _pt.tree_map(lambda g: g.mul_(step_size), updates)
        else:
            updates = _pt.tree_map(lambda g: g.mul(step_size), updates)
        return updates, state

    return GradientTransformation(init_fn, update_fn)


def chain(*transforms):
    def init_fn(params):
        return tuple(t.init(params) for t in transforms)

    def update_fn(updates, state, *, params=None, inplace=True):
        new_state = []
        for t, s in zip(transforms, state):
            updates, ns = t.update(updates, s, params=params, inplace=inplace)
            new_state.append(ns)
        return updates, tuple(new_state)

    return GradientTransformation(init_fn, update_fn)


def adam(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, *, eps_root=0.0,
         moment_requires_grad=False, maximize=False, use_accelerated_op=False):
    if weight_decay != 0.0 or maximize:
        raise NotImplementedError("oracle shim restates adam with weight_decay=0, maximize=False only")
    b1, b2 = betas
    return chain(_scale_by_adam(b1, b2, eps, eps_root), _scale(-lr))


def sgd(lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, *, maximize=False):
    if momentum != 0.0 or weight_decay != 0.0 or nesterov or maximize:
        raise NotImplementedError("oracle shim restates plain sgd only")
    return chain(_scale(-lr))


def apply_updates(params, updates, *, inplace=True):
    if inplace:
        _pt.tree_map(lambda p, u: p.add_(u), params, updates)
        return params
    return _pt.tree_map(lambda p, u: p.add(u), params, updates)
