"""Turn a Python right-hand side callback into a device functor.

The reference accepts ANY Python function ``f(ozone_vars, *args, **kwargs)`` and lets csdl record and
differentiate it (``ozone/func_wrapper.py:90-107``), including with respect to tensors the function
captures from its closure (``test_3_large/utils.py:18,30``).  Here the callback is executed ONCE on
symbolic inputs: every scalar that flows through it is an expression node, arrays of scalars are
numpy object arrays wrapped in ``TVar`` (so indexing, reshaping, broadcasting, ``+ - * / **`` and the
functions in ``ozone_b200.ops`` work the way a csdl/numpy user expects).  From the recorded
expressions this module

  * differentiates symbolically (Jacobian for the implicit stage solve, VJPs w.r.t. states, dynamic
    parameters and captured tensors for the reverse sweep, output VJPs),
  * detects trailing quadrature states (zero Jacobian columns -> ``DQ``),
  * emits a CUDA functor with the contract of ``csrc/oz_rhs.cuh`` and compiles it with nvcc against the
    same kernel templates as the built-in right-hand sides (one stage count / dtype / approach per
    build, a few seconds), caching the shared object under ``build/plugins/<hash>.so``,
  * registers it with ``oz_register_rhs_plugin``.

The callback must be node-wise (row i of every output depends on row i of the inputs only) -- true of
every right-hand side in the reference -- and its per-trajectory state must be small enough for one
thread (packed size <= 32): large independent systems are expressed as a batch of trajectories.
"""
from __future__ import annotations

import hashlib
import math
import os
import subprocess
from typing import Optional

import numpy as np
import torch

from . import _native as nat

_HERE = os.path.dirname(os.path.abspath(__file__))
PLUGIN_DIR = os.environ.get('OZONE_B200_PLUGIN_DIR') or os.path.join(os.path.dirname(_HERE), 'build', 'plugins')
MAX_STATE = 32
MAX_STATIC = 64
_REGISTERED: dict = {}     # plugin path -> rhs_id (the library keeps a fixed table of registrations)


# ================================================================================================
# expression DAG (hash-consed)
class Expr:
    __slots__ = ('op', 'args', 'val', 'idx', '_h')
    __array_priority__ = 1000
    _table: dict = {}

    def __new__(cls, op, args=(), val=None, idx=None):
        key = (op, tuple(id(a) for a in args), val, idx)
        hit = cls._table.get(key)
        if hit is not None:
            return hit
        self = object.__new__(cls)
        self.op, self.args, self.val, self.idx = op, tuple(args), val, idx
        self._h = key
        cls._table[key] = self
        return self

    # ---- constructors with light algebraic simplification
    @staticmethod
    def const(v):
        return Expr('const', val=float(v))

    @staticmethod
    def wrap(v):
        if isinstance(v, Expr):
            return v
        if isinstance(v, (int, float, np.floating, np.integer)):
            return Expr.const(v)
        raise TypeError(f'cannot use {type(v).__name__} in a traced right-hand side')

    def is_const(self, v=None):
        return self.op == 'const' and (v is None or self.val == v)

    def __add__(self, o):
        o = Expr.wrap(o)
        if self.is_const() and o.is_const():
            return Expr.const(self.val + o.val)
        if self.is_const(0.0):
            return o
        if o.is_const(0.0):
            return self
        return Expr('add', (self, o))

    __radd__ = lambda self, o: Expr.wrap(o).__add__(self)

    def __sub__(self, o):
        o = Expr.wrap(o)
        if self.is_const() and o.is_const():
            return Expr.const(self.val - o.val)
        if o.is_const(0.0):
            return self
        if self.is_const(0.0):
            return -o
        return Expr('sub', (self, o))

    __rsub__ = lambda self, o: Expr.wrap(o).__sub__(self)

    def __mul__(self, o):
        o = Expr.wrap(o)
        if self.is_const() and o.is_const():
            return Expr.const(self.val * o.val)
        if self.is_const(0.0) or o.is_const(0.0):
            return Expr.const(0.0)
        if self.is_const(1.0):
            return o
        if o.is_const(1.0):
            return self
        if o.is_const():
            self, o = o, self                     # constants first, so that nested constant factors fold
        if self.is_const() and o.op == 'mul' and o.args[0].is_const():
            return Expr.const(self.val * o.args[0].val) * o.args[1]
        return Expr('mul', (self, o))

    __rmul__ = lambda self, o: Expr.wrap(o).__mul__(self)

    def __truediv__(self, o):
        o = Expr.wrap(o)
        if self.is_const() and o.is_const():
            return Expr.const(self.val / o.val)
        if self.is_const(0.0):
            return self
        if o.is_const(1.0):
            return self
        return Expr('div', (self, o))

    __rtruediv__ = lambda self, o: Expr.wrap(o).__truediv__(self)

    def __neg__(self):
        if self.is_const():
            return Expr.const(-self.val)
        if self.op == 'neg':
            return self.args[0]
        return Expr('neg', (self,))

    def __pos__(self):
        return self

    def __pow__(self, e):
        if isinstance(e, Expr) and e.is_const():
            e = e.val
        if isinstance(e, (int, float, np.floating, np.integer)):
            e = float(e)
            if e == 0.0:
                return Expr.const(1.0)
            if e == 1.0:
                return self
            if self.is_const():
                return Expr.const(self.val ** e)
            return Expr('powc', (self,), val=e)
        return Expr('exp', (Expr.wrap(e) * Expr('log', (self,)),))

    def __rpow__(self, base):          # const ** expr
        return Expr('exp', (self * Expr.const(math.log(float(base))),))

    def _fn(self, name, pyf):
        if self.is_const():
            return Expr.const(pyf(self.val))
        return Expr(name, (self,))

    # numpy ufuncs on object arrays call these method names
    def exp(self): return self._fn('exp', math.exp)
    def log(self): return self._fn('log', math.log)
    def sin(self): return self._fn('sin', math.sin)
    def cos(self): return self._fn('cos', math.cos)
    def tan(self): return self._fn('tan', math.tan)
    def tanh(self): return self._fn('tanh', math.tanh)
    def sqrt(self): return self._fn('sqrt', math.sqrt)
    def arctan(self): return self._fn('atan', math.atan)

    def __abs__(self):
        return self._fn('abs', abs)

    def __bool__(self):
        raise TypeError('the truth value of a traced quantity is not known while tracing: data-dependent branches '
                        'cannot be compiled; use smooth expressions (ozone_b200.ops) instead')

    def __float__(self):
        if self.is_const():
            return self.val
        raise TypeError('a traced quantity has no numeric value while tracing')


def leaf(kind, i):
    return Expr(kind, idx=int(i))


def diff(e: Expr, var: Expr, memo: dict) -> Expr:
    """d e / d var (var is a leaf), memoised per (e, var)."""
    if e is var:
        return Expr.const(1.0)
    key = (id(e), id(var))
    if key in memo:
        return memo[key]
    op, a = e.op, e.args
    if not a:
        r = Expr.const(0.0)
    elif op == 'add':
        r = diff(a[0], var, memo) + diff(a[1], var, memo)
    elif op == 'sub':
        r = diff(a[0], var, memo) - diff(a[1], var, memo)
    elif op == 'mul':
        r = diff(a[0], var, memo) * a[1] + a[0] * diff(a[1], var, memo)
    elif op == 'div':
        r = diff(a[0], var, memo) / a[1] - e * diff(a[1], var, memo) / a[1]
    elif op == 'neg':
        r = -diff(a[0], var, memo)
    else:
        da = diff(a[0], var, memo)
        if da.is_const(0.0):
            r = da
        elif op == 'powc':
            r = Expr.const(e.val) * (a[0] ** (e.val - 1.0)) * da
        elif op == 'exp':
            r = e * da
        elif op == 'log':
            r = da / a[0]
        elif op == 'sin':
            r = a[0].cos() * da
        elif op == 'cos':
            r = -(a[0].sin()) * da
        elif op == 'tan':
            r = (Expr.const(1.0) + e * e) * da
        elif op == 'tanh':
            r = (Expr.const(1.0) - e * e) * da
        elif op == 'sqrt':
            r = da / (Expr.const(2.0) * e)
        elif op == 'atan':
            r = da / (Expr.const(1.0) + a[0] * a[0])
        elif op == 'abs':
            r = (a[0] / e) * da
        else:
            raise NotImplementedError(op)
    memo[key] = r
    return r


# ================================================================================================
# arrays of expressions
class TVar:
    """An array of traced scalars with the small csdl/numpy surface right-hand sides use."""
    __array_priority__ = 1000

    def __array_ufunc__(self, ufunc, method, *inputs, **kwargs):
        # np.exp(x), np.add(x, 1.0), ndarray + TVar ...: run the ufunc on the underlying object arrays,
        # where numpy falls back to the Expr operators / methods (exp, sin, sqrt, ...)
        if method != '__call__' or kwargs.get('out') is not None:
            return NotImplemented
        r = ufunc(*[_arr(x) for x in inputs])
        return TVar(r if isinstance(r, np.ndarray) else np.asarray(r, dtype=object).reshape(()))

    def __init__(self, a):
        self.a = a if isinstance(a, np.ndarray) and a.dtype == object else np.asarray(a, dtype=object)

    # ---- shape
    shape = property(lambda self: self.a.shape)
    ndim = property(lambda self: self.a.ndim)
    size = property(lambda self: self.a.size)

    def numel(self):
        return self.a.size

    def dim(self):
        return self.a.ndim

    def __len__(self):
        return len(self.a)

    def __getitem__(self, k):
        r = self.a[k]
        return TVar(r) if isinstance(r, np.ndarray) else TVar(np.asarray(r, dtype=object).reshape(()))

    def reshape(self, *shape):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list)):
            shape = tuple(shape[0])
        return TVar(self.a.reshape(shape))

    def flatten(self):
        return TVar(self.a.reshape(-1))

    @property
    def T(self):
        return TVar(self.a.T)

    def expand(self, shape, action=None):
        """csdl's ``x.expand(shape, action='i->ij')``: result[dst] = x[src], new letters are broadcast.
        Without an action: plain numpy broadcasting (a scalar to any shape)."""
        shape = tuple(shape)
        if action is None:
            return TVar(np.broadcast_to(self.a, shape).copy())
        src, dst = action.replace(' ', '').split('->')
        a = self.a.reshape(self.a.shape) if self.a.ndim == len(src) else self.a.reshape((1,) * (len(src) - self.a.ndim) + self.a.shape)
        order = [src.index(c) for c in dst if c in src]           # source axes in the order they appear in dst
        a = np.transpose(a, order)
        it = iter(a.shape)
        a = a.reshape(tuple(next(it) if c in src else 1 for c in dst))
        return TVar(np.broadcast_to(a, shape).copy())

    def set(self, key, value):
        """csdl's functional ``x.set(csdl.slice[...], v)``: returns a copy with the slice replaced."""
        out = self.a.copy()
        out[key] = _arr(value)
        return TVar(out)

    def sum(self, axis=None, dim=None, axes=None):
        ax = axis if axis is not None else (dim if dim is not None else axes)
        if isinstance(ax, list):
            ax = tuple(ax)
        r = self.a.sum(axis=ax)
        return TVar(r if isinstance(r, np.ndarray) else np.asarray(r, dtype=object).reshape(()))

    # ---- arithmetic
    def _bin(self, o, fn):
        return TVar(fn(self.a, _arr(o)))

    __add__ = lambda s, o: s._bin(o, lambda a, b: a + b)
    __radd__ = lambda s, o: s._bin(o, lambda a, b: b + a)
    __sub__ = lambda s, o: s._bin(o, lambda a, b: a - b)
    __rsub__ = lambda s, o: s._bin(o, lambda a, b: b - a)
    __mul__ = lambda s, o: s._bin(o, lambda a, b: a * b)
    __rmul__ = lambda s, o: s._bin(o, lambda a, b: b * a)
    __truediv__ = lambda s, o: s._bin(o, lambda a, b: a / b)
    __rtruediv__ = lambda s, o: s._bin(o, lambda a, b: b / a)
    __neg__ = lambda s: TVar(-s.a)
    __pos__ = lambda s: s

    def __pow__(self, e):
        if isinstance(e, (int, float, np.floating, np.integer)):
            return TVar(np.vectorize(lambda x: x ** e, otypes=[object])(self.a))
        return self._bin(e, lambda a, b: a ** b)

    def __rpow__(self, base):
        return TVar(np.vectorize(lambda x: x.__rpow__(base), otypes=[object])(self.a))

    def __matmul__(self, o):
        return TVar(self.a @ _arr(o))

    def __rmatmul__(self, o):
        return TVar(_arr(o) @ self.a)

    def map(self, name):
        return TVar(np.vectorize(lambda x: getattr(Expr.wrap(x), name)(), otypes=[object])(self.a))

    def __abs__(self):
        return TVar(np.vectorize(abs, otypes=[object])(self.a))

    # torch.Tensor (op) TVar lands here: the tensor is a closure-captured static parameter
    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        name = getattr(func, '__name__', '')
        table = {'add': '__radd__', 'sub': '__rsub__', 'mul': '__rmul__', 'div': '__rtruediv__',
                 'true_divide': '__rtruediv__', 'pow': '__rpow__', 'matmul': '__rmatmul__',
                 '__add__': '__radd__', '__sub__': '__rsub__', '__mul__': '__rmul__', '__truediv__': '__rtruediv__',
                 '__pow__': '__rpow__', '__matmul__': '__rmatmul__', '__radd__': '__add__', '__rmul__': '__mul__',
                 '__rsub__': '__sub__', '__rtruediv__': '__truediv__'}
        if name in table and len(args) == 2 and isinstance(args[1], TVar):
            return getattr(args[1], table[name])(args[0])
        if name in table and len(args) == 2 and isinstance(args[0], TVar):
            return getattr(args[0], table[name].replace('__r', '__'))(args[1])
        raise TypeError(f'torch.{name} is not available on traced quantities: use operators or ozone_b200.ops')

    def __repr__(self):
        return f'TVar(shape={self.shape})'


_ACTIVE: list = []        # tracer stack: captured tensors are registered with the innermost one


def _arr(o):
    """operand -> numpy array (object or float) usable against an object array of Expr."""
    if isinstance(o, TVar):
        return o.a
    if isinstance(o, torch.Tensor):
        if not _ACTIVE:
            raise RuntimeError('tensor met outside of a trace')
        return _ACTIVE[-1].capture(o).a
    if isinstance(o, Expr):
        return np.asarray(o, dtype=object).reshape(())
    return np.asarray(o, dtype=np.float64) if not isinstance(o, np.ndarray) or o.dtype != object else o


def build_and_register_plugin(src: str, flags, stem: str, what: str) -> int:
    """nvcc the generated translation unit into ``build/plugins/<stem>_<hash>.so`` (cached: the hash covers the source, the
    flags and every kernel template header) and register it with the library; returns the rhs_id."""
    csrc = os.path.join(_HERE, 'csrc')
    stamp = hashlib.sha256()
    stamp.update(src.encode())
    stamp.update(' '.join(flags).encode())
    for fn in sorted(os.listdir(csrc)):                      # kernel templates are part of the plugin
        if fn.endswith(('.cuh', '.h')):
            stamp.update(open(os.path.join(csrc, fn), 'rb').read())
    tag = stamp.hexdigest()[:20]
    os.makedirs(PLUGIN_DIR, exist_ok=True)
    so = os.path.join(PLUGIN_DIR, f'{stem}_{tag}.so')
    if not os.path.exists(so):
        # every rank under torchrun may build the same plugin at once: private file names, atomic publish
        cu = os.path.join(PLUGIN_DIR, f'{stem}_{tag}.{os.getpid()}.cu')
        tmp = f'{so}.{os.getpid()}.tmp'
        open(cu, 'w').write(src)
        cmd = ['nvcc', '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-std=c++17', '-lineinfo',
               '--expt-relaxed-constexpr', '-split-compile', '0', '-Xcompiler', '-fPIC', '-Xcompiler', '-fvisibility=hidden',
               '-shared', f'-I{csrc}', *flags, cu, '-o', tmp, '-lcudart_static', '-lrt', '-lpthread', '-ldl']
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f'nvcc failed on {what}:\n{res.stderr[-4000:]}')
        os.replace(tmp, so)
    rid = _REGISTERED.get(so)                 # one registration per shared object and process
    if rid is None:
        rid = nat.lib.oz_register_rhs_plugin(so.encode())
        if rid < 0:
            raise nat.OzError(f'oz_register_rhs_plugin: {nat.lib.oz_last_error().decode()}')
        _REGISTERED[so] = rid
    return rid


# ================================================================================================
class Tracer:
    """Records one call of a callback on symbolic inputs."""

    def __init__(self, batch_size=1):
        self.static = []            # [(tensor, offset, point shape)]
        self.n_static = 0
        self.batch_size = batch_size

    def capture(self, t: torch.Tensor) -> TVar:
        """A tensor met inside the callback (closure-captured or passed through set_function) becomes a
        static parameter.  In a batched problem a trailing axis of length N is the trajectory axis."""
        per_traj = self.batch_size > 1 and t.dim() >= 1 and t.shape[-1] == self.batch_size
        view_shape = tuple(t.shape[:-1]) if per_traj else tuple(t.shape)
        for tt, off, shp in self.static:
            if tt is t:
                break
        else:
            off, shp = self.n_static, (view_shape if view_shape else (1,))
            n = int(math.prod(shp))
            if self.n_static + n > MAX_STATIC:
                raise ValueError(f'a traced right-hand side may capture at most {MAX_STATIC} static scalars; '
                                 f'pass large data as dynamic parameters or states')
            self.static.append((t, off, shp))
            self.n_static += n
        a = np.empty(int(math.prod(shp)), dtype=object)
        for i in range(a.size):
            a[i] = leaf('th', off + i)
        return TVar(a.reshape(view_shape))


TRACE_NODES = 3       # the callback is traced on three nodes: shapes are checked like the reference does with
                      # num_nodes > 1, and node 0's outputs must not depend on node 1's inputs (node-wise)


def _sym(kind, offset, shape):
    n = int(math.prod(shape)) if shape else 1
    a = np.empty((TRACE_NODES, n), dtype=object)
    for node in range(TRACE_NODES):
        for i in range(n):
            a[node, i] = leaf(kind if node == 0 else f'{kind}_other{node}', offset + i)
    return a.reshape((TRACE_NODES,) + (tuple(shape) if shape else (1,)))   # scalars are (num_nodes, 1), like csdl


def _leaves(exprs):
    seen, out, stack = set(), set(), list(exprs)
    while stack:
        e = stack.pop()
        if id(e) in seen:
            continue
        seen.add(id(e))
        if e.args:
            stack.extend(e.args)
        elif e.op != 'const':
            out.add(e.op)
    return out


def _norm_shape(s):
    s = tuple(s)
    return () if s == (1,) else s


class TracedRHS:
    """What ``FuncWrapper`` turns a Python callback into: same role as a ``rhs.DeviceRHS`` handle."""
    kind = 'batched'
    config_parameters = ()

    def __init__(self, f, args, kwargs, state_shapes: dict, param_shapes: dict, batch_size: int = 1):
        from .func_wrapper import ODEVars
        self.state_decl = [(n, tuple(s)) for n, s in state_shapes.items()]
        self.param_decl = [(n, tuple(s)) for n, s in param_shapes.items()]
        tr = Tracer(batch_size)
        states, params, off = {}, {}, 0
        for n, s in self.state_decl:
            states[n] = TVar(_sym('y', off, s))
            off += int(math.prod(s)) if s else 1
        self.D = off
        off = 0
        for n, s in self.param_decl:
            params[n] = TVar(_sym('p', off, s))
            off += int(math.prod(s)) if s else 1
        self.P = off
        if self.D > MAX_STATE:
            raise ValueError(f'the packed state of one trajectory has {self.D} components; traced right-hand sides '
                             f'support up to {MAX_STATE} per trajectory. Express independent systems as a batch '
                             f'(trailing axis = trajectories, ODEProblem(..., batched=True)).')
        ov = ODEVars(states, params, TRACE_NODES)
        _ACTIVE.append(tr)
        try:
            targs = [tr.capture(a) if isinstance(a, torch.Tensor) else a for a in args]
            tkw = {k: (tr.capture(v) if isinstance(v, torch.Tensor) else v) for k, v in kwargs.items()}
            f(ov, *targs, **tkw)
            ov.post_process()
        finally:
            _ACTIVE.pop()
        self.tracer = tr
        self.S = tr.n_static
        flat = lambda v: [Expr.wrap(x) for x in _arr(v)[0].reshape(-1)]          # node 0
        self.f_exprs = []
        for n, s in self.state_decl:
            self.f_exprs += flat(ov.d_states[n])
        self.prof_decl = [(n, _norm_shape(v.shape[1:])) for n, v in ov.profile_outputs.items()]
        self.field_decl = [(n, _norm_shape(v.shape[1:])) for n, v in ov.field_outputs.items()]
        self.prof_exprs = [e for n, _ in self.prof_decl for e in flat(ov.profile_outputs[n])]
        self.field_exprs = [e for n, _ in self.field_decl for e in flat(ov.field_outputs[n])]
        outputs = list(ov.d_states.values()) + list(ov.profile_outputs.values()) + list(ov.field_outputs.values())
        for node in range(TRACE_NODES):         # row `node` of every output may only read row `node` of the inputs
            kinds = _leaves([Expr.wrap(x) for v in outputs for x in _arr(v)[node].reshape(-1)])
            own = {'th'} | ({'y', 'p'} if node == 0 else {f'y_other{node}', f'p_other{node}'})
            if kinds - own:
                raise ValueError('the ODE function is not node-wise: a row of its outputs depends on a different row of '
                                 'its inputs; every row (time point / stage) must be computed from the same row')
        self._analyse()
        self._plugins = {}

    # ---- the DeviceRHS surface used by collections.Packing / integrators
    def names(self, which):
        return {'states': self.state_decl, 'dynamic_parameters': self.param_decl,
                'static_parameters': [(f'_captured{i}', _norm_shape(shp)) for i, (_, _, shp) in enumerate(self.tracer.static)],
                'profile_outputs': self.prof_decl, 'field_outputs': self.field_decl}[which]

    static_parameters = property(lambda self: self.names('static_parameters'))
    profile_outputs = property(lambda self: self.prof_decl)
    field_outputs = property(lambda self: self.field_decl)

    def bind_static(self, args, kwargs):
        return {f'_captured{i}': t for i, (t, _, _) in enumerate(self.tracer.static)}

    def rhs_cfg(self, static_values, states):
        return ()

    # ---- symbolic analysis
    def _analyse(self):
        D, P, S = self.D, self.P, self.S
        memo = {}
        y = [leaf('y', i) for i in range(D)]
        p = [leaf('p', i) for i in range(P)]
        th = [leaf('th', i) for i in range(S)]
        self.J = [[diff(self.f_exprs[i], y[j], memo) for j in range(D)] for i in range(D)]
        v = [leaf('v', i) for i in range(D)]

        def vjp(outs, cots, wrt):
            res = []
            for w in wrt:
                acc = Expr.const(0.0)
                for e, c in zip(outs, cots):
                    acc = acc + c * diff(e, w, memo)
                res.append(acc)
            return res
        self.vjp_y, self.vjp_p, self.vjp_th = vjp(self.f_exprs, v, y), vjp(self.f_exprs, v, p), vjp(self.f_exprs, v, th)
        vp = [leaf('vp', i) for i in range(len(self.prof_exprs))]
        vf = [leaf('vf', i) for i in range(len(self.field_exprs))]
        outs, cots = self.prof_exprs + self.field_exprs, vp + vf
        self.ovjp_y, self.ovjp_p, self.ovjp_th = vjp(outs, cots, y), vjp(outs, cots, p), vjp(outs, cots, th)
        dq = 0
        for j in range(D - 1, -1, -1):          # trailing states nothing depends on
            if all(self.J[i][j].is_const(0.0) for i in range(D)):
                dq += 1
            else:
                break
        self.DQ = dq if dq < D else D - 1

    # ---- code generation
    def cuda_source(self) -> str:
        def block(assignments, accumulate=False):
            """assignments: [(lhs, Expr)] -> C statements with shared temporaries."""
            lines, names, count = [], {}, [0]

            def emit(e: Expr) -> str:
                if id(e) in names:
                    return names[id(e)]
                op = e.op
                if op == 'const':
                    s = f'T({e.val!r})' if math.isfinite(e.val) else ('T(CUDART_INF)' if e.val > 0 else 'T(-CUDART_INF)')
                elif op in ('y', 'p', 'th', 'v', 'vp', 'vf'):
                    s = f'{op}[{e.idx}]'
                else:
                    a = [emit(x) for x in e.args]
                    if op == 'add': s = f'({a[0]} + {a[1]})'
                    elif op == 'sub': s = f'({a[0]} - {a[1]})'
                    elif op == 'mul': s = f'({a[0]} * {a[1]})'
                    elif op == 'div': s = f'({a[0]} / {a[1]})'
                    elif op == 'neg': s = f'(-{a[0]})'
                    elif op == 'powc':
                        k = e.val
                        if k == 2.0: s = f'({a[0]} * {a[0]})'
                        elif k == 3.0: s = f'({a[0]} * {a[0]} * {a[0]})'
                        elif k == -1.0: s = f'(T(1) / {a[0]})'
                        elif k == 0.5: s = f'sqrt({a[0]})'
                        else: s = f'pow({a[0]}, T({k!r}))'
                    elif op == 'abs': s = f'fabs({a[0]})'
                    else: s = f'{op}({a[0]})'
                    name = f't{count[0]}'
                    count[0] += 1
                    lines.append(f'        const T {name} = {s};')
                    s = name
                names[id(e)] = s
                return s
            for lhs, e in assignments:
                if accumulate and e.is_const(0.0):
                    continue
                lines.append(f'        {lhs} {"+=" if accumulate else "="} {emit(e)};')
            return '\n'.join(lines)

        D, P, S = self.D, self.P, self.S
        NP, NF = len(self.prof_exprs), len(self.field_exprs)
        f_body = block([(f'dy[{i}]', e) for i, e in enumerate(self.f_exprs)])
        j_body = block([(f'J[{i * D + j}]', self.J[i][j]) for i in range(D) for j in range(D)])
        v_body = block([(f'yb[{j}]', e) for j, e in enumerate(self.vjp_y)] + [(f'pb[{j}]', e) for j, e in enumerate(self.vjp_p)]
                       + [(f'tb[{j}]', e) for j, e in enumerate(self.vjp_th)], accumulate=True)
        o_body = block([(f'prof[{i}]', e) for i, e in enumerate(self.prof_exprs)]
                       + [(f'field[{i}]', e) for i, e in enumerate(self.field_exprs)])
        ov_body = block([(f'yb[{j}]', e) for j, e in enumerate(self.ovjp_y)] + [(f'pb[{j}]', e) for j, e in enumerate(self.ovjp_p)]
                        + [(f'tb[{j}]', e) for j, e in enumerate(self.ovjp_th)], accumulate=True)
        jp = ', '.join('0' if self.J[i][j].is_const(0.0) else '1' for i in range(D) for j in range(D))
        memo = {}
        pl = [leaf('p', j) for j in range(P)]
        jpp = ', '.join('0' if diff(self.f_exprs[i], pl[j], memo).is_const(0.0) else '1' for i in range(D) for j in range(P))
        jpp_line = (f'    static constexpr unsigned char JPP[{D * P}] = {{{jpp}}};   // the same for df_r/dp_c\n' if P else '')
        return f'''// generated by ozone_b200/trace.py from a Python right-hand side callback
#include "oz_common.cuh"
namespace oz {{
struct RhsTraced {{
    static constexpr int ID = OZ_RHS_PLUGIN_BASE, D = {D}, P = {P}, S = {S}, NPROF = {NP}, NFIELD = {NF};
    static constexpr int DQ = {self.DQ};
    static constexpr unsigned char JP[{D * D}] = {{{jp}}};   // which df_r/dy_c are not identically zero (oz_sparse_solve.cuh)
{jpp_line}    template <typename T>
    OZ_D static void f(const T* y, const T* p, const T* th, T* dy) {{
{f_body}
    }}
    template <typename T>
    OZ_D static void jac(const T* y, const T* p, const T* th, T* J) {{
{j_body}
    }}
    template <typename T>
    OZ_D static void vjp(const T* y, const T* p, const T* th, const T* v, T* yb, T* pb, T* tb) {{
{v_body}
    }}
    template <typename T>
    OZ_D static void out(const T* y, const T* p, const T* th, T* prof, T* field) {{
{o_body}
    }}
    template <typename T>
    OZ_D static void out_vjp(const T* y, const T* p, const T* th, const T* vp, const T* vf, T* yb, T* pb, T* tb) {{
{ov_body}
    }}
}};
}}  // namespace oz
#include "oz_plugin_impl.cuh"
'''

    # ---- compile + register: one build per (stages, implicit, dtype, kind)
    def plugin_id(self, num_stages: int, implicit: bool, fp64: bool, kind: int) -> int:
        key = (num_stages, bool(implicit), bool(fp64), kind)
        if key in self._plugins:
            return self._plugins[key]
        flags = [f'-DOZ_ONLY_S={num_stages}', f'-DOZ_ONLY_IMPL={int(implicit)}', f'-DOZ_ONLY_F64={int(fp64)}',
                 f'-DOZ_ONLY_KIND={kind}']
        rid = build_and_register_plugin(self.cuda_source(), flags, 'oz_rhs', 'the traced right-hand side')
        self._plugins[key] = rid
        return rid

    def __repr__(self):
        return f'<traced RHS D={self.D} P={self.P} S={self.S} DQ={self.DQ}>'
