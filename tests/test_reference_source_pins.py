"""Pins of this repository's DATA on the reference's own source, wherever that source can be imported.

The reference as a whole cannot run here (csdl_alpha is not installable), but two parts of it can:
  * ``ozone/glms/**`` is numpy only: all 22 registered tableaux are compared bit for bit with
    ``oracle.tableaux`` and ``ozone_b200.glms``;
  * the right-hand side callbacks of the examples are ordinary Python functions of ``ozone_vars``: they are
    traced by ``ozone_b200.trace`` exactly like a user callback, and (a) the generated device functor must be
    identical to the one generated from the re-typed copy the GPU tests run (``tests/reference_callbacks.py``),
    (b) their values and Jacobians must equal the oracle's restatement (``oracle/rhs.py``) at random points;
  * the PDE example's callback (``pde_control.py:11-81``, written with csdl functions over the whole grid) goes through
    ``ozone_b200.trace_grid`` with ``csdl`` bound to ``ozone_b200.ops``: same two checks.
``/root/reference`` exists only in the build container: where it is absent the live comparisons are skipped and the
committed fixture ``tests/golden/reference_tableaux.json`` (written by ``tools/make_reference_goldens.py`` from the
reference's module) carries the tableau pin.
"""
import json
import math
import os

import numpy as np
import pytest
import torch

from oracle import rhs as oracle_rhs
from oracle.tableaux import tableau
from ozone_b200 import trace
from ozone_b200.glms import get_integration_method
from tests import reference_callbacks as RC
from tests import reference_source as RS
from tests.test_trace_symbolic import evaluate

needs_reference = pytest.mark.skipif(not RS.AVAILABLE, reason='/root/reference is not present on this machine')
GOLDEN = os.path.join(os.path.dirname(__file__), 'golden', 'reference_tableaux.json')


def _same_bits(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return a.shape == b.shape and np.array_equal(a.view(np.int64), b.view(np.int64))


@needs_reference
def test_all_registered_tableaux_equal_the_reference_module_bit_for_bit():
    with RS.reference_glms() as G:
        names = sorted(G.method_classes)
        assert len(names) == 22                                            # ozone/glms/GLMs.py:45-68
        for name in names:
            ref = G.get_integration_method(name)                           # GLMs.py:31-42
            A, B = np.asarray(ref.A, dtype=np.float64), np.asarray(ref.B, dtype=np.float64).reshape(-1)
            oa, ob, oc, oexpl = tableau(name)
            assert _same_bits(A, oa) and _same_bits(B, ob) and bool(ref.explicit) == bool(oexpl), name
            assert _same_bits(np.asarray(ref.C, dtype=np.float64).reshape(-1), oc), name
            ours = get_integration_method(name)
            assert _same_bits(A, ours.A) and _same_bits(B, np.asarray(ours.B).reshape(-1)), name
            assert bool(ours.explicit) == bool(ref.explicit) and ours.num_stages == ref.num_stages, name
            assert _same_bits(np.asarray(ref.C, dtype=np.float64), ours.C), name
            assert _same_bits(np.asarray(ref.U, dtype=np.float64).reshape(-1), np.ones(len(A))), name   # U = 1, V = 1
            assert _same_bits(np.asarray(ref.V, dtype=np.float64).reshape(-1), np.ones(1)), name


def test_tableaux_equal_the_fixture_written_from_the_reference_module():
    gold = json.load(open(GOLDEN))
    assert len(gold) == 22
    for name, g in gold.items():
        A = np.array([[float.fromhex(x) for x in row] for row in g['A']])
        B = np.array([float.fromhex(x) for x in g['B']])
        oa, ob, _, oexpl = tableau(name)
        ours = get_integration_method(name)
        assert _same_bits(A, oa) and _same_bits(B, ob) and g['explicit'] == bool(oexpl), name
        assert _same_bits(A, ours.A) and _same_bits(B, np.asarray(ours.B).reshape(-1)), name
        assert g['explicit'] == bool(ours.explicit), name


def _trace(problem, f):
    states, params = RC.DECLARATIONS[problem]
    args = (torch.tensor([0.5], dtype=torch.float64),) if problem == 'lotka_volterra' else ()
    return trace.TracedRHS(f, args, {}, states, params)


@needs_reference
@pytest.mark.parametrize('problem', sorted(RC.CALLBACKS))
def test_reference_callback_and_its_test_copy_generate_the_same_device_functor(problem):
    ref = _trace(problem, RS.reference_callback(problem))
    ours = _trace(problem, RC.CALLBACKS[problem])
    assert ref.cuda_source() == ours.cuda_source()
    assert (ref.D, ref.P, ref.S, ref.DQ) == (ours.D, ours.P, ours.S, ours.DQ)


@needs_reference
@pytest.mark.parametrize('problem', sorted(RC.CALLBACKS))
def test_oracle_right_hand_sides_equal_the_traced_reference_callbacks(problem):
    """f, df/dy, and the outputs of the oracle's restatement against the reference's own function, evaluated through
    the tracer's expression DAG (plain Python floats) at random points."""
    t = _trace(problem, RS.reference_callback(problem))
    orc = oracle_rhs.REGISTRY[problem]()
    rng = np.random.default_rng(7)
    for _ in range(5):
        y = rng.uniform(0.5, 2.0, size=t.D)
        p = rng.uniform(-1.0, 1.0, size=t.P)
        th = rng.uniform(0.4, 0.6, size=t.S)
        env = {('y', i): float(v) for i, v in enumerate(y)}
        env.update({('p', i): float(v) for i, v in enumerate(p)})
        env.update({('th', i): float(v) for i, v in enumerate(th)})
        f_ref = np.array([evaluate(e, env) for e in t.f_exprs])
        J_ref = np.array([[evaluate(e, env) for e in row] for row in t.J])
        Y, Pm, Th = y[:, None], p[:, None], (th[:, None] if t.S else None)
        f_or = np.asarray(orc.f(Y, Pm, Th))[:, 0]
        J_or = np.asarray(orc.jac(Y, Pm, Th))[..., 0]
        assert np.allclose(f_or, f_ref, rtol=1e-14, atol=1e-15), problem
        assert np.allclose(J_or, J_ref, rtol=1e-14, atol=1e-15), problem
        if t.prof_exprs or t.field_exprs:
            prof, fld = orc.out(Y, Pm, Th)
            assert np.allclose(np.asarray(prof)[:, 0], [evaluate(e, env) for e in t.prof_exprs], rtol=1e-14)
            assert np.allclose(np.asarray(fld)[:, 0], [evaluate(e, env) for e in t.field_exprs], rtol=1e-14)


# ---------------------------------------------------------------- the grid callback (pde_control.py:11-81)
PDE_KW = dict(zero_bc=0, pi_bc=0, beta_t=16.0, beta_u=2.0, gamma=2.0)


def _trace_grid(f, nx):
    from ozone_b200.trace_grid import TracedGridRHS
    return TracedGridRHS(f, (), dict(dx=np.pi / nx, nx=nx, **PDE_KW), 'x', nx, ['u'])


@needs_reference
def test_reference_pde_callback_and_the_test_callback_are_the_same_grid_model():
    """The reference's own ``ode_function`` (csdl bound to ozone_b200.ops) and the callback the GPU tests run: same coefficient
    field, and the same cell expression, diagonal, neighbour weights, control weight and field-output integrand at random points."""
    from ozone_b200 import ops
    from tests import grid_callbacks as G
    nx = 90
    ref = _trace_grid(RS.reference_callback('pde_control', {'csdl': ops}), nx)
    ours = _trace_grid(G.pde_control, nx)
    assert np.array_equal(ref.tracer.coefficient_table(), ours.tracer.coefficient_table())
    rng = np.random.default_rng(5)
    for _ in range(10):
        xl, x, xr, c, u = rng.uniform(0.05, 1.2, 5)
        env = {('gx', -1): xl, ('gx', 0): x, ('gx', 1): xr, ('gc', 0): c, ('p', 0): u}
        for name in ('f', 'dg', 'wr_of_left', 'wl_of_right', 'q', 'dq', 'r'):
            a, b = evaluate(getattr(ref, name), env), evaluate(getattr(ours, name), env)
            assert a == pytest.approx(b, rel=1e-13, abs=1e-13), name
        assert evaluate(ref.du[0], env) == pytest.approx(evaluate(ours.du[0], env), rel=1e-14)
        assert evaluate(ref.dr[0], env) == pytest.approx(evaluate(ours.dr[0], env), rel=1e-14)


@needs_reference
def test_oracle_pde_equals_the_traced_reference_callback_cell_by_cell():
    from ozone_b200 import ops
    nx = 40
    t = _trace_grid(RS.reference_callback('pde_control', {'csdl': ops}), nx)
    orc = oracle_rhs.DiffusionReaction(nx)
    tab = t.tracer.coefficient_table()
    rng = np.random.default_rng(11)
    for _ in range(3):
        x, u = rng.uniform(0.1, 1.0, size=nx), float(rng.uniform(-0.6, 0.0))
        f_or = orc.f(x[:, None], np.array([[u]]), None)[:, 0]
        J_or = float(orc.out(x[:, None], np.array([[u]]), None)[1][0, 0])
        xe = np.concatenate(([0.0], x, [0.0]))                     # zero ghost cells (:32-35)
        f_tr, q_sum = np.empty(nx), 0.0
        for k in range(nx):
            env = {('gx', -1): xe[k], ('gx', 0): xe[k + 1], ('gx', 1): xe[k + 2], ('p', 0): u}
            env.update({('gc', q): tab[q, k] for q in range(tab.shape[0])})
            f_tr[k] = evaluate(t.f, env)
            q_sum += evaluate(t.q, env)
        assert np.allclose(f_tr, f_or, rtol=1e-13, atol=1e-12)
        assert q_sum + evaluate(t.r, {('p', 0): u}) == pytest.approx(J_or, rel=1e-13)
