"""The tracer's expression DAG and its symbolic derivatives (ozone_b200/trace.py), checked on the CPU with an
evaluator written here: every operation a callback can record, differentiated and compared with central
differences, plus the simplifications that keep generated functors small."""
import math

import numpy as np
import pytest

from ozone_b200.trace import Expr, diff, leaf

FN = {'exp': math.exp, 'log': math.log, 'sin': math.sin, 'cos': math.cos, 'tan': math.tan, 'tanh': math.tanh,
      'sqrt': math.sqrt, 'atan': math.atan, 'abs': abs}


def evaluate(e, env, memo=None):
    memo = {} if memo is None else memo
    if id(e) in memo:
        return memo[id(e)]
    a = [evaluate(x, env, memo) for x in e.args]
    if e.op == 'const':
        v = e.val
    elif not e.args:
        v = env[(e.op, e.idx)]
    elif e.op == 'add':
        v = a[0] + a[1]
    elif e.op == 'sub':
        v = a[0] - a[1]
    elif e.op == 'mul':
        v = a[0] * a[1]
    elif e.op == 'div':
        v = a[0] / a[1]
    elif e.op == 'neg':
        v = -a[0]
    elif e.op == 'powc':
        v = a[0] ** e.val
    else:
        v = FN[e.op](a[0])
    memo[id(e)] = v
    return v


Y0, Y1, P0 = leaf('y', 0), leaf('y', 1), leaf('p', 0)
CASES = {
    'polynomial': lambda y0, y1, p: (1.0 - y1 * y1) * y0 - y1 + p,
    'quotient': lambda y0, y1, p: (y0 + 2.0 * p) / (1.5 + y1 * y1) - y0 / 3.0,
    'powers': lambda y0, y1, p: y0 ** 3 - (y1 + 2.0) ** 0.5 + p ** -2.0 + 2.0 ** y0 + y0 ** y1,
    'transcendental': lambda y0, y1, p: (y0 * p).exp() * y1.sin() + (y0 + 3.0).log() * y1.cos() + y0.tanh() - p.arctan(),
    'sqrt_abs_tan': lambda y0, y1, p: (y0 * y0 + y1 * y1 + 1.0).sqrt() + abs(y0 - 2.0 * y1) + (0.3 * p).tan(),
    'arrhenius': lambda y0, y1, p: 16.0 * ((-2.0 / (1.0 + y0)).exp() - math.exp(-2.0)) + 2.0 * (p - y0),   # pde_control.py:45-52
}


@pytest.mark.parametrize('name', sorted(CASES))
def test_symbolic_derivatives_match_central_differences(name):
    e = CASES[name](Y0, Y1, P0)
    point = {('y', 0): 0.7, ('y', 1): -0.4, ('p', 0): 1.3}
    for var in (Y0, Y1, P0):
        d = evaluate(diff(e, var, {}), point)
        key, eps = (var.op, var.idx), 1e-6
        hi, lo = dict(point), dict(point)
        hi[key] += eps
        lo[key] -= eps
        fd = (evaluate(e, hi) - evaluate(e, lo)) / (2 * eps)
        assert d == pytest.approx(fd, rel=2e-8, abs=1e-9), (name, key)


def test_hash_consing_and_simplifications():
    a = Y0 * Y1 + P0
    assert a is Y0 * Y1 + P0                                    # one node per distinct expression
    assert (Y0 + 0.0) is Y0 and (Y0 * 1.0) is Y0 and (Y0 - 0.0) is Y0 and (Y0 / 1.0) is Y0
    assert (Y0 * 0.0).is_const(0.0) and (0.0 * Y0).is_const(0.0)
    assert (-(-Y0)) is Y0 and (Y0 ** 1) is Y0 and (Y0 ** 0).is_const(1.0)
    assert (Expr.const(2.0) * Expr.const(3.0)).is_const(6.0)    # constants fold while tracing
    # a state the expression does not touch differentiates to the constant 0: this is how zero Jacobian columns
    # (quadrature states) are detected
    assert diff(Y0 * Y0 + P0 * P0, Y1, {}).is_const(0.0)


def test_data_dependent_branches_are_refused():
    with pytest.raises(TypeError):
        bool(Y0 - 1.0)
    with pytest.raises(TypeError):
        float(Y0)
    with pytest.raises(TypeError):
        Expr.wrap('text')


def test_second_derivative_of_the_van_der_pol_cost_is_constant():
    J = Y0 * Y0 + Y1 * Y1 + P0 * P0                             # van_der_pol/main_model.py:12
    assert evaluate(diff(diff(J, Y0, {}), Y0, {}), {}) == 2.0
    assert diff(diff(J, Y0, {}), Y1, {}).is_const(0.0)


def test_generated_functor_compiles_and_registers_without_a_gpu():
    """Trace the reference's Lotka-Volterra callback (lotka_volterra.py:51-62), emit the CUDA functor, build the
    plugin with nvcc (cross-compiles for sm_100a here) and register it through the C ABI.  No kernel is launched."""
    import torch
    from ozone_b200 import trace
    c = torch.tensor(0.5, dtype=torch.float64)

    def ode_function(v):
        a, x, y = v.dynamic_parameters['a'], v.states['x'], v.states['y']
        v.d_states['y'] = a * y - 0.5 * y * x
        v.d_states['x'] = 2.0 * x * y - c * x
        v.field_outputs['field_x'] = x
        v.profile_outputs['x_plus_y'] = x + y
    t = trace.TracedRHS(ode_function, (), {}, {'y': (), 'x': ()}, {'a': ()})
    assert (t.D, t.P, t.S, t.DQ) == (2, 1, 1, 0)            # the captured tensor c became a static parameter
    src = t.cuda_source()
    for piece in ('struct RhsTraced', 'static void f(', 'static void jac(', 'static void vjp(', 'static void out_vjp('):
        assert piece in src, piece
    rid = t.plugin_id(num_stages=4, implicit=False, fp64=True, kind=0)
    assert rid >= 1000                                       # OZ_RHS_PLUGIN_BASE in include/ozone_b200.h
    assert t.plugin_id(num_stages=4, implicit=False, fp64=True, kind=0) == rid       # cached: one registration
