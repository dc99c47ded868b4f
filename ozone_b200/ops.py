"""Functions a right-hand side callback may apply to its variables (the csdl functions the reference's
callbacks use: exp, sin, sum, expand, reshape, concatenate, ...).  They work on traced quantities
(``trace.TVar``, while a Python callback is being compiled to a device functor) and on torch tensors."""
import numpy as np
import torch

from .trace import TVar, _arr
from . import trace_grid as _grid
from .trace_grid import GridVar


def _unary(name, tfn):
    def fn(x):
        if isinstance(x, (TVar, GridVar)):
            return x.map(name)
        return tfn(torch.as_tensor(x))
    fn.__name__ = name
    return fn


exp = _unary('exp', torch.exp)
log = _unary('log', torch.log)
sin = _unary('sin', torch.sin)
cos = _unary('cos', torch.cos)
tan = _unary('tan', torch.tan)
tanh = _unary('tanh', torch.tanh)
sqrt = _unary('sqrt', torch.sqrt)
arctan = _unary('arctan', torch.atan)


def sum(x, axes=None, axis=None):
    ax = axes if axes is not None else axis
    if isinstance(x, GridVar):
        return x.sum()
    if isinstance(x, TVar):
        return x.sum(axis=tuple(ax) if isinstance(ax, (list, tuple)) else ax)
    return torch.sum(x) if ax is None else torch.sum(x, dim=ax)


def expand(x, shape, action=None):
    tr = _grid.active_grid_tracer()
    if tr is not None and not isinstance(x, GridVar) and int(np.prod(shape)) == tr.nx and tr.nx > 1:
        # a scalar spread over the cells of a traced grid right-hand side
        v = x.a.reshape(-1) if isinstance(x, TVar) else np.asarray(x, dtype=object).reshape(-1)
        if v.size != 1:
            raise ValueError('only scalars can be expanded over the grid')
        return GridVar(tr, _grid.Expr.wrap(v[0]), shape)
    if isinstance(x, TVar):
        return x.expand(shape, action)
    from .trace import _ACTIVE
    if isinstance(x, torch.Tensor):
        if not _ACTIVE:                                     # eager evaluation on plain tensors
            if action is None:
                return x.expand(tuple(shape)).clone()
            src, dst = action.replace(' ', '').split('->')
            view = x.permute([src.index(c) for c in dst if c in src])
            view = view.reshape(tuple(view.shape[[c for c in dst if c in src].index(c)] if c in src else 1 for c in dst))
            return view.expand(tuple(shape)).clone()
        return TVar(_arr(x)).expand(shape, action)          # captured tensor used inside a trace
    return TVar(np.asarray(x, dtype=object)).expand(shape, action)


def reshape(x, shape):
    return x.reshape(shape)


def concatenate(xs, axis=0):
    if any(isinstance(x, GridVar) for x in xs):
        return _grid.concatenate(xs, axis)
    if any(isinstance(x, TVar) for x in xs):
        return TVar(np.concatenate([_arr(x) for x in xs], axis=axis))
    ref = next(x for x in xs if isinstance(x, torch.Tensor))       # eager evaluation: numpy pieces join the tensors
    return torch.cat([x if isinstance(x, torch.Tensor) else torch.as_tensor(np.asarray(x), dtype=ref.dtype, device=ref.device)
                      for x in xs], dim=axis)


def norm(x):
    if isinstance(x, TVar):
        return (x * x).sum().map('sqrt')
    return torch.linalg.norm(x)
