"""Turn a Python right-hand side written over a 1-D grid into a device grid model (grid form).

The reference's PDE example hands ``set_function`` an ordinary callback whose state is the whole grid and which builds
the stencil from shifted slices (``ozone/paper_examples/pde_control/pde_control.py:11-81``)::

    x_vec0 = csdl.concatenate((zeros, x_vec[:-1, :]))        # left neighbour, zero ghost cell
    x_vec1 = csdl.concatenate((x_vec[1:, :], zeros))         # right neighbour
    dxdz2 = (x_vec0 + (-2.0) * x_vec + x_vec1) / dx**2
    ...  + beta_t * np.exp(1) ** (-gamma * (1 + x_vec) ** (-1))  +  expand(u, (nx, 1)) * (beta_u * heaviside_array)

Such a callback is run ONCE here on a symbolic grid: a ``GridVar`` is "one scalar expression per cell", written relative
to the cell (``x[k-1], x[k], x[k+1]``, per-cell coefficient fields, the shared dynamic parameter).  Slicing the cell axis
shifts that expression, concatenation with zeros is the zero ghost cell, numpy arrays with one entry per cell become
coefficient fields that the kernels read from HBM, ``sum`` over the grid becomes the integrand of a field output.
From the recorded expression this module differentiates symbolically and emits a grid model with the contract of
``csrc/oz_pde_kernels.cuh`` -- the same fused multi-step forward / reverse kernels as the built-in diffusion-reaction
system, ``exp`` and reciprocals through the same fast device routines -- and compiles it into a plugin
(``csrc/oz_grid_plugin_impl.cuh``), cached under ``build/plugins``.

What the kernels can express, and what is refused with an error that says so:
  * one state (the grid), up to four scalar dynamic parameters, explicit methods, TimeMarching / TimeMarchingCheckpoints;
  * a three-point stencil with constant (zero or non-zero) or reflecting boundary values whose NEIGHBOUR terms are linear, with weights that are constants or
    coefficient fields (variable diffusivity, upwind advection); the cell's own value may enter in any smooth way;
  * the dynamic parameters enter linearly (their weights may be coefficient fields);
  * at most one field output of the form  const * sum_cells q(x, coefficients) + r(parameter); no profile outputs;
  * captured data must be floats or numpy arrays (they are constants of the generated code / coefficient fields), not
    tensors to differentiate.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from .trace import Expr, TVar, leaf, diff, _ACTIVE, build_and_register_plugin

_UNARY = {'exp': math.exp, 'log': math.log, 'sin': math.sin, 'cos': math.cos, 'tan': math.tan, 'tanh': math.tanh,
          'sqrt': math.sqrt, 'atan': math.atan, 'abs': abs}
_UFUNC_BIN = {'add': ('__add__', '__radd__'), 'subtract': ('__sub__', '__rsub__'), 'multiply': ('__mul__', '__rmul__'),
              'divide': ('__truediv__', '__rtruediv__'), 'true_divide': ('__truediv__', '__rtruediv__'),
              'power': ('__pow__', '__rpow__')}
_UFUNC_UN = {'exp': 'exp', 'log': 'log', 'sin': 'sin', 'cos': 'cos', 'tan': 'tan', 'tanh': 'tanh', 'sqrt': 'sqrt',
             'arctan': 'atan', 'absolute': 'abs', 'negative': 'neg'}


def _rebuild(e: Expr, args):
    """The node `e` with new arguments, through the simplifying constructors."""
    op = e.op
    if op == 'add':
        return args[0] + args[1]
    if op == 'sub':
        return args[0] - args[1]
    if op == 'mul':
        return args[0] * args[1]
    if op == 'div':
        return args[0] / args[1]
    if op == 'neg':
        return -args[0]
    if op == 'powc':
        return args[0] ** e.val
    return args[0]._fn(op, _UNARY[op])


def substitute(e: Expr, fn, memo: dict) -> Expr:
    """Replace leaves: fn(leaf) -> Expr (or None to keep it)."""
    k = id(e)
    if k in memo:
        return memo[k]
    if not e.args:
        r = e
        if e.op != 'const':
            got = fn(e)
            if got is not None:
                r = got
    else:
        r = _rebuild(e, [substitute(a, fn, memo) for a in e.args])
    memo[k] = r
    return r


def leaves_of(e: Expr):
    seen, out, stack = set(), [], [e]
    while stack:
        x = stack.pop()
        if id(x) in seen:
            continue
        seen.add(id(x))
        if x.args:
            stack.extend(x.args)
        elif x.op != 'const':
            out.append(x)
    return out


class GridTracer:
    """Bookkeeping of one trace: coefficient fields and grid sums met by the callback."""

    def __init__(self, nx: int):
        self.nx = int(nx)
        self.bases = []        # distinct per-cell arrays (float64, nx)
        self._base_key = {}
        self.coefs = []        # [(base index, shift)]: the coefficient fields the generated code reads
        self._coef_key = {}
        self.sums = []         # integrands of sum-over-the-grid nodes (leaf 'gs' k)

    def capture(self, t):
        raise TypeError('a grid right-hand side cannot capture tensors: pass floats (constants of the generated code) or '
                        'numpy arrays with one entry per cell (coefficient fields)')

    def coef_leaf(self, base: int, shift: int) -> Expr:
        key = (base, shift)
        if key not in self._coef_key:
            self._coef_key[key] = len(self.coefs)
            self.coefs.append(key)
        return leaf('gc', self._coef_key[key])

    def field(self, values) -> Expr:
        a = np.ascontiguousarray(np.asarray(values, dtype=np.float64).reshape(-1))
        key = a.tobytes()
        if key not in self._base_key:
            self._base_key[key] = len(self.bases)
            self.bases.append(a)
        return self.coef_leaf(self._base_key[key], 0)

    def shift(self, e: Expr, d: int) -> Expr:
        """The expression of element k + d written for element k."""
        if d == 0:
            return e

        def move(l):
            if l.op == 'gx':
                return leaf('gx', l.idx + d)
            if l.op == 'gc':
                base, sh = self.coefs[l.idx]
                return self.coef_leaf(base, sh + d)
            return None
        return substitute(e, move, {})

    def coefficient_table(self) -> np.ndarray:
        """[K, nx]: row q is coefficient field q at every cell (shifted copies are zero beyond the ends)."""
        nx = self.nx
        out = np.zeros((len(self.coefs), nx))
        for q, (base, sh) in enumerate(self.coefs):
            src = self.bases[base]
            lo, hi = max(0, -sh), min(nx, nx - sh)          # cells k with 0 <= k + sh < nx
            if hi > lo:
                out[q, lo:hi] = src[lo + sh:hi + sh]
        return out


def _zero_when_state_is_zero(e: Expr) -> bool:
    r = substitute(e, lambda l: Expr.const(0.0) if l.op == 'gx' else None, {})
    return r.is_const(0.0)


class GridVar:
    """One traced scalar per cell.  `expr` is written for "this element": leaf gx[s] is the state s cells to the right of it,
    gc[q] coefficient field q at it.  `shape` has at most one axis longer than 1 (the cells)."""
    __array_priority__ = 2000

    def __init__(self, tracer: GridTracer, expr: Expr, shape):
        self.tr, self.expr, self.shape = tracer, expr, tuple(int(s) for s in shape)
        self.n = int(math.prod(self.shape))
        if sum(1 for s in self.shape if s != 1) > 1:
            raise ValueError(f'a grid quantity has one cell axis; shape {self.shape} has several axes longer than 1')

    ndim = property(lambda self: len(self.shape))
    size = property(lambda self: self.n)

    def numel(self):
        return self.n

    def __len__(self):
        return self.shape[0]

    # ---- shape
    def reshape(self, *shape):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list)):
            shape = tuple(shape[0])
        shape = list(shape)
        if -1 in shape:
            known = int(math.prod(s for s in shape if s != -1))
            shape[shape.index(-1)] = self.n // max(known, 1)
        if int(math.prod(shape)) != self.n:
            raise ValueError(f'cannot reshape a grid quantity of {self.n} cells into shape {tuple(shape)}')
        return GridVar(self.tr, self.expr, shape)

    def flatten(self):
        return self.reshape(self.n)

    @property
    def T(self):
        return GridVar(self.tr, self.expr, self.shape[::-1])

    def __getitem__(self, key):
        key = key if isinstance(key, tuple) else (key,)
        if any(k is Ellipsis for k in key):
            raise TypeError('"..." is not supported on a grid quantity')
        key = key + (slice(None),) * (len(self.shape) - len(key))
        if len(key) != len(self.shape):
            raise IndexError(f'too many indices for a grid quantity of shape {self.shape}')
        shape, expr = [], self.expr
        for k, s in zip(key, self.shape):
            if isinstance(k, (int, np.integer)):
                if s != 1:
                    raise TypeError('single cells of a grid state cannot be addressed in a grid right-hand side: the same '
                                    'expression has to hold for every cell (use slices, and coefficient arrays for what '
                                    'differs from cell to cell)')
                if int(k) not in (0, -1):
                    raise IndexError(f'index {k} out of range for an axis of length 1')
                continue                                      # the axis is dropped
            if not isinstance(k, slice) or k.step not in (None, 1):
                raise TypeError(f'unsupported index {k!r} on a grid quantity (slices with step 1 only)')
            lo, hi, _ = k.indices(s)
            if hi <= lo:
                raise ValueError('empty slice of a grid quantity')
            shape.append(hi - lo)
            if s != 1 and lo:
                expr = self.tr.shift(expr, lo)             # new element k' = old element k' + lo
        return GridVar(self.tr, expr, shape)

    # ---- arithmetic
    def _operand(self, o):
        """-> (Expr for this element, shape)."""
        if isinstance(o, GridVar):
            if o.n != self.n and o.n != 1 and self.n != 1:
                raise ValueError(f'grid quantities of {self.n} and {o.n} cells cannot be combined')
            return o.expr, o.shape
        if isinstance(o, torch.Tensor):
            self.tr.capture(o)
        if isinstance(o, TVar):
            if o.size != 1:
                raise ValueError('only scalars (one value per time point) can be combined with a grid quantity')
            return Expr.wrap(o.a.reshape(-1)[0]), ()
        if isinstance(o, Expr):
            return o, ()
        a = np.asarray(o, dtype=np.float64)
        if a.size == 1:
            return Expr.const(float(a.reshape(-1)[0])), ()
        if a.size != self.n or self.n != self.tr.nx:
            raise ValueError(f'an array of {a.size} values cannot be combined with a grid quantity of {self.n} cells '
                             f'(coefficient arrays have one entry per cell of the whole grid, {self.tr.nx})')
        if sum(1 for s in a.shape if s != 1) > 1:
            raise ValueError(f'coefficient array of shape {a.shape}: expected one axis of {self.n} cells')
        return self.tr.field(a), a.shape

    def _shape_with(self, oshape):
        try:
            shp = np.broadcast_shapes(self.shape, tuple(oshape))
        except ValueError:
            raise ValueError(f'shapes {self.shape} and {tuple(oshape)} of two grid quantities do not match') from None
        if int(math.prod(shp)) != max(self.n, int(math.prod(oshape)) if oshape else 1):
            raise ValueError(f'shapes {self.shape} and {tuple(oshape)} would broadcast to a matrix; reshape one of them')
        return shp

    def _bin(self, o, fn):
        e, shp = self._operand(o)
        return GridVar(self.tr, fn(self.expr, e), self._shape_with(shp))

    __add__ = lambda s, o: s._bin(o, lambda a, b: a + b)
    __radd__ = lambda s, o: s._bin(o, lambda a, b: b + a)
    __sub__ = lambda s, o: s._bin(o, lambda a, b: a - b)
    __rsub__ = lambda s, o: s._bin(o, lambda a, b: b - a)
    __mul__ = lambda s, o: s._bin(o, lambda a, b: a * b)
    __rmul__ = lambda s, o: s._bin(o, lambda a, b: b * a)
    __truediv__ = lambda s, o: s._bin(o, lambda a, b: a / b)
    __rtruediv__ = lambda s, o: s._bin(o, lambda a, b: b / a)
    __neg__ = lambda s: GridVar(s.tr, -s.expr, s.shape)
    __pos__ = lambda s: s

    def __pow__(self, e):
        if isinstance(e, (int, float, np.floating, np.integer)):
            return GridVar(self.tr, self.expr ** e, self.shape)
        return self._bin(e, lambda a, b: a ** b)

    def __rpow__(self, base):
        return GridVar(self.tr, self.expr.__rpow__(float(base)), self.shape)

    def map(self, name):
        name = {'arctan': 'atan'}.get(name, name)
        if name == 'neg':
            return -self
        return GridVar(self.tr, self.expr._fn(name, _UNARY[name]), self.shape)

    def __abs__(self):
        return self.map('abs')

    exp = lambda s: s.map('exp')
    log = lambda s: s.map('log')
    sin = lambda s: s.map('sin')
    cos = lambda s: s.map('cos')
    tan = lambda s: s.map('tan')
    tanh = lambda s: s.map('tanh')
    sqrt = lambda s: s.map('sqrt')

    def __array_ufunc__(self, ufunc, method, *inputs, **kwargs):
        if method != '__call__' or kwargs.get('out') is not None:
            return NotImplemented
        name = ufunc.__name__
        if name in _UFUNC_BIN and len(inputs) == 2:
            fwd, rev = _UFUNC_BIN[name]
            a, b = inputs
            return getattr(a, fwd)(b) if isinstance(a, GridVar) else getattr(b, rev)(a)
        if name in _UFUNC_UN and len(inputs) == 1:
            return self.map(_UFUNC_UN[name])
        return NotImplemented

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        raise TypeError('a grid right-hand side cannot be combined with tensors: pass floats or numpy arrays')

    # ---- reductions
    def sum(self, axis=None, dim=None, axes=None):
        if self.n != self.tr.nx:
            raise ValueError('only sums over the whole grid are supported in a grid right-hand side')
        for l in leaves_of(self.expr):
            if l.op == 'gx' and l.idx != 0:
                raise NotImplementedError('a sum over the grid may only use the cell\'s own value, not its neighbours')
            if l.op not in ('gx', 'gc'):
                raise NotImplementedError('a sum over the grid may only depend on the state and on coefficient arrays; '
                                          'multiply by parameters outside of the sum')
        self.tr.sums.append(self.expr)
        return TVar(np.asarray(leaf('gs', len(self.tr.sums) - 1), dtype=object).reshape(()))

    def __repr__(self):
        return f'GridVar(shape={self.shape})'


def concatenate(pieces, axis=0):
    """Ghost cells: ``concatenate((zeros, x[:-1]))`` / ``concatenate((x[1:], zeros))``.

    The longest grid piece is the shifted state; the kernels evaluate it with ZERO ghost cells, so what is concatenated to it
    is added on top as coefficient fields that only the padded cells see: constants (Dirichlet data) as their values, other
    grid pieces -- ``x[:1]`` in front for a reflecting boundary -- as (0/1 indicator) x (that piece's expression)."""
    grid = [p for p in pieces if isinstance(p, GridVar)]
    if not grid:
        raise TypeError('internal: grid concatenation without a grid quantity')
    g = max(grid, key=lambda p: p.n)                      # the shifted state; everything else is boundary data
    tr = g.tr
    offsets, n = [], 0
    for p in pieces:
        offsets.append(n)
        if isinstance(p, GridVar):
            n += p.n
        else:
            a = p.a if isinstance(p, TVar) else p
            a = np.asarray(a.detach().cpu() if isinstance(a, torch.Tensor) else a)
            try:
                n += a.astype(np.float64).size
            except (TypeError, ValueError):
                raise NotImplementedError('only constants or slices of the state can be concatenated to a grid quantity '
                                          '(boundary values); a traced scalar cannot be a ghost cell') from None
    front = next(o for p, o in zip(pieces, offsets) if p is g)
    back = n - front - g.n
    expr = tr.shift(g.expr, -front)
    if front or back:
        if n != tr.nx:
            raise NotImplementedError('padding must restore the full grid length')
        if not _zero_when_state_is_zero(expr):
            raise NotImplementedError('a padded slice must vanish where the state vanishes (it stands for ghost cells): '
                                      'shift the state itself, then apply nonlinear functions')
        for l in leaves_of(expr):
            if l.op == 'gx' and ((front and l.idx > -front) or (back and l.idx < back)):
                raise NotImplementedError('the padded cells would not be ghost cells: a slice x[a:] goes with padding at the END, '
                                          'x[:-a] with padding at the FRONT')
        pad = np.zeros(n)
        for p, o in zip(pieces, offsets):
            if p is g:
                continue
            if isinstance(p, GridVar):                    # e.g. x[:1] in front: the ghost cell mirrors the boundary cell
                ind = np.zeros(n)
                ind[o:o + p.n] = 1.0
                expr = expr + tr.field(ind) * tr.shift(p.expr, -o)
            else:
                a = p.a if isinstance(p, TVar) else p
                a = np.asarray(a.detach().cpu() if isinstance(a, torch.Tensor) else a).astype(np.float64).reshape(-1)
                pad[o:o + a.size] = a
        if np.any(pad != 0.0):            # Dirichlet data: zero ghost cell + a constant that only the boundary cells see
            expr = expr + tr.field(pad)
    elif len(pieces) != 1:
        raise NotImplementedError('empty pieces in a grid concatenation')
    shape = list(g.shape)
    cell_axis = next((i for i, s in enumerate(shape) if s != 1), axis if axis < len(shape) else 0) if shape else 0
    if not shape:
        shape = [n]
    else:
        shape[cell_axis] = n
    return GridVar(tr, expr, shape)


def active_grid_tracer():
    return _ACTIVE[-1] if _ACTIVE and isinstance(_ACTIVE[-1], GridTracer) else None


# ================================================================================================
class TracedGridRHS:
    """What ``FuncWrapper`` turns a Python callback over a grid into: same role as ``rhs.diffusion_reaction``."""
    kind = 'grid'
    config_parameters = ()
    COEF_NAME = '_grid_coefficients'
    MAX_PARAMS = 4                 # kPdeMaxP in csrc/oz_pde_kernels.cuh

    def __init__(self, f, args, kwargs, state_name: str, nx: int, param_names):
        from .func_wrapper import ODEVars
        if len(param_names) > self.MAX_PARAMS:
            raise NotImplementedError(f'a grid right-hand side takes at most {self.MAX_PARAMS} (scalar) dynamic parameters; got '
                                      f'{list(param_names)}')
        self.state_name, self.nx = state_name, int(nx)
        self.param_names = list(param_names)
        tr = GridTracer(nx)
        states = {state_name: GridVar(tr, leaf('gx', 0), (1, nx))}
        params = {n: TVar(np.asarray([[leaf('p', i)]], dtype=object)) for i, n in enumerate(self.param_names)}
        ov = ODEVars(states, params, 1)
        _ACTIVE.append(tr)
        try:
            f(ov, *args, **kwargs)
            ov.post_process()
        finally:
            _ACTIVE.pop()
        self.tracer = tr
        if ov.profile_outputs:
            raise NotImplementedError('profile outputs are not available for a grid right-hand side')
        if len(ov.field_outputs) > 1:
            raise NotImplementedError('a grid right-hand side supports one field output')
        dy = ov.d_states[state_name]
        if not isinstance(dy, GridVar):
            raise ValueError(f"the derivative of grid state '{state_name}' must be computed from the state")
        self.f = dy.expr
        self.field_decl = [(n, ()) for n in ov.field_outputs]
        self.J = None
        for n, v in ov.field_outputs.items():
            if isinstance(v, GridVar) or v.size != 1:
                raise ValueError(f"field output '{n}' must be one value per time point (sum over the grid)")
            self.J = Expr.wrap(v.a.reshape(-1)[0])
        self._analyse()
        self._plugins = {}

    # ---- the DeviceRHS surface used by collections.Packing / integrators
    def names(self, which):
        K = len(self.tracer.coefs)
        return {'states': [(self.state_name, None)], 'dynamic_parameters': [(n, ()) for n in self.param_names],
                'static_parameters': [(self.COEF_NAME, (K * self.nx,))] if K else [],
                'profile_outputs': [], 'field_outputs': self.field_decl}[which]

    static_parameters = property(lambda self: self.names('static_parameters'))
    profile_outputs = ()
    field_outputs = property(lambda self: self.field_decl)

    def bind_static(self, args, kwargs):
        if not self.tracer.coefs:
            return {}
        return {self.COEF_NAME: torch.from_numpy(self.tracer.coefficient_table().reshape(-1))}

    def rhs_cfg(self, static_values, states):
        return ()

    def spec_info(self, d_override):
        return dict(d=int(d_override), P=len(self.param_names), S=len(self.tracer.coefs) * self.nx, nprof=0,
                    nfield=len(self.field_decl), kind=1)

    # ---- symbolic analysis
    def _analyse(self):
        tr, memo = self.tracer, {}
        kinds = {}
        for l in leaves_of(self.f):
            kinds.setdefault(l.op, set()).add(l.idx)
        if 'gs' in kinds:
            raise NotImplementedError('the state derivative may not depend on a sum over the grid (non-local coupling)')
        if kinds.get('gx', set()) - {-1, 0, 1}:
            raise NotImplementedError(f'the kernels implement a three-point stencil; this right-hand side reaches cells '
                                      f'{sorted(kinds["gx"])} away')
        xl, xc, xr = leaf('gx', -1), leaf('gx', 0), leaf('gx', 1)
        wl, wr = diff(self.f, xl, memo), diff(self.f, xr, memo)
        for w in (wl, wr):
            if any(l.op != 'gc' for l in leaves_of(w)):
                raise NotImplementedError('neighbour cells must enter linearly, with constant or per-cell weights '
                                          '(the cell\'s own value may enter in any smooth way)')
        # (J^T G)_j = (weight of the right neighbour in cell j-1) G_{j-1} + dg_j G_j + (weight of the left neighbour in cell j+1) G_{j+1}
        self.wr_of_left, self.wl_of_right = tr.shift(wr, -1), tr.shift(wl, 1)
        self.dg = diff(self.f, xc, memo)
        self.du = []
        for q, name in enumerate(self.param_names):
            d = diff(self.f, leaf('p', q), memo)
            if any(l.op != 'gc' for l in leaves_of(d)):
                raise NotImplementedError(f"dynamic parameter '{name}' must enter the state derivative linearly, with a constant "
                                          f'or per-cell weight')
            self.du.append(d)
        # field output: J = sum_k a_k sum_cells q_k + r(u)
        self.q, self.r = Expr.const(0.0), Expr.const(0.0)
        if self.J is not None:
            for k, integrand in enumerate(tr.sums):
                a = diff(self.J, leaf('gs', k), memo)
                if not a.is_const():
                    raise NotImplementedError('a field output must be  const * sum_cells(...) + r(parameter)')
                self.q = self.q + a * integrand
            self.r = substitute(self.J, lambda l: Expr.const(0.0) if l.op == 'gs' else None, {})
            if any(l.op != 'p' for l in leaves_of(self.r)):
                raise NotImplementedError('a field output must be  const * sum_cells(...) + r(parameter)')
        self.dq = diff(self.q, xc, memo)
        self.dr = [diff(self.r, leaf('p', q), memo) for q in range(len(self.param_names))]

    # ---- code generation
    @staticmethod
    def _block(assignments, fast: bool):
        """[(lhs, Expr)] -> C statements with shared temporaries.  fast: exp and reciprocals through the grid kernels' own
        device routines (pde_exp needs the CTA's table `etab`)."""
        lines, names, count = [], {}, [0]

        def emit(e: Expr) -> str:
            if id(e) in names:
                return names[id(e)]
            op = e.op
            if op == 'const':
                s = f'T({e.val!r})' if math.isfinite(e.val) else ('T(CUDART_INF)' if e.val > 0 else 'T(-CUDART_INF)')
            elif op == 'gx':
                s = {-1: 'xl', 0: 'x', 1: 'xr'}[e.idx]
            elif op == 'gc':
                s = f'c[{e.idx}]'
            elif op == 'p':
                s = f'u[{e.idx}]'
            elif op == 'v':                       # arguments of jt()
                s = ('gl', 'g', 'gr', 'dg')[e.idx]
            else:
                a = [emit(x) for x in e.args]
                if op == 'add': s = f'({a[0]} + {a[1]})'
                elif op == 'sub': s = f'({a[0]} - {a[1]})'
                elif op == 'mul': s = f'({a[0]} * {a[1]})'
                elif op == 'div':
                    if e.args[1].is_const() and e.args[1].val != 0.0 and math.isfinite(1.0 / e.args[1].val):
                        s = f'({a[0]} * T({1.0 / e.args[1].val!r}))'
                    else:
                        s = f'({a[0]} * pde_rcp({a[1]}))'
                elif op == 'neg': s = f'(-{a[0]})'
                elif op == 'powc':
                    k = e.val
                    if k == 2.0: s = f'({a[0]} * {a[0]})'
                    elif k == 3.0: s = f'({a[0]} * {a[0]} * {a[0]})'
                    elif k == -1.0: s = f'pde_rcp({a[0]})'
                    elif k == -2.0: s = f'pde_rcp({a[0]} * {a[0]})'
                    elif k == 0.5: s = f'sqrt({a[0]})'
                    else: s = f'pow({a[0]}, T({k!r}))'
                elif op == 'abs': s = f'fabs({a[0]})'
                elif op == 'exp': s = f'pde_exp({a[0]}, etab)' if fast else f'exp({a[0]})'
                else: s = f'{op}({a[0]})'
                name = f't{count[0]}'
                count[0] += 1
                lines.append(f'        const T {name} = {s};')
                s = name
            names[id(e)] = s
            return s
        for lhs, e in assignments:
            lines.append(f'        {lhs} {emit(e)};')
        return '\n'.join(lines)

    def cuda_source(self) -> str:
        K = len(self.tracer.coefs)
        both = self._block([('dg =', self.dg), ('return', self.f)], fast=True)
        only = self._block([('return', self.f)], fast=True)
        jt = self._block([('return', self.wr_of_left * leaf('v', 0) + self.wl_of_right * leaf('v', 2) + leaf('v', 1) * leaf('v', 3))],
                         fast=False)
        per_q = lambda exprs, prefix: '\n'.join(
            f'        if (q == {q}) {{\n' + self._block([(prefix, e)], fast=False) + '\n        }' for q, e in enumerate(exprs))
        du = per_q(self.du, 'return') + '\n        return T(0);'
        q = self._block([('return', self.q)], fast=False)
        dq = self._block([('return', self.dq)], fast=False)
        r = self._block([('return own *', self.r)], fast=False)
        dr = per_q(self.dr, 'return fb *') + '\n        return T(0);'
        return f'''// generated by ozone_b200/trace_grid.py from a Python right-hand side callback over a 1-D grid
#include "oz_pde_kernels.cuh"
namespace oz {{
template <typename T_>
struct GridTraced {{
    using T = T_;
    static constexpr int NCOEF = {K}, NFIELD = {len(self.field_decl)};
    static constexpr int P = {len(self.param_names)};
    static constexpr bool WINDOW = false;
    void configure(const oz_problem*, long long, T* field_scale, T* fb_scale) {{
        *field_scale = T(1);
        *fb_scale = T(1);
    }}
    OZ_D T stage_u(T ulo, T du, T cs) const {{ return ulo + cs * du; }}
    template <bool WITH_DIAG>
    OZ_D T f(T xl, T x, T xr, T, const T* c, const T* u, const T* etab, T& dg) const {{
        if constexpr (WITH_DIAG) {{
{both}
        }} else {{
{only}
        }}
    }}
    OZ_D T jt(T gl, T g, T gr, T dg, const T* c) const {{
{jt}
    }}
    OZ_D T du(T, const T* c, int q) const {{
{du}
    }}
    OZ_D T du_scale() const {{ return T(1); }}
    OZ_D T q(T x, const T* c) const {{
{q}
    }}
    OZ_D T dq(T x, const T* c) const {{
{dq}
    }}
    OZ_D T ucost(T own, const T* u) const {{
{r}
    }}
    OZ_D T ducost(T fb, const T* u, int q) const {{
{dr}
    }}
}};
}}  // namespace oz
#include "oz_grid_plugin_impl.cuh"
'''

    # ---- compile + register: one build per (stages, dtype)
    def plugin_id(self, num_stages: int, implicit: bool, fp64: bool, kind: int) -> int:
        if implicit:
            raise ValueError('a grid right-hand side supports explicit methods only (an implicit step would be an '
                             f'{self.nx} x {self.nx} Newton system per step)')
        if kind != 0:
            raise ValueError('a grid right-hand side runs under TimeMarching / TimeMarchingCheckpoints only')
        key = (num_stages, bool(fp64))
        if key in self._plugins:
            return self._plugins[key]
        flags = [f'-DOZ_ONLY_S={num_stages}', f'-DOZ_ONLY_F64={int(fp64)}']
        rid = build_and_register_plugin(self.cuda_source(), flags, 'oz_grid', 'the traced grid right-hand side')
        self._plugins[key] = rid
        return rid

    def __repr__(self):
        return f'<traced grid RHS nx={self.nx} coefficient fields={len(self.tracer.coefs)}>'
