"""The reference's Collocation tests (test_1_simple/test_solution.py:65-81, test_2_general/test_solution.py:65-81):
the harness hands the declared design variables / constraints to SLSQP with a dummy objective
(ozone/tests/run_ode.py:24-33, 49-58) and checks the converged states against the same goldens as the other
approaches (derivative checks are disabled there).  Here scipy's SLSQP plays modopt's part."""
import numpy as np
import pytest
import torch

import ozone_b200 as ozone
from ozone_b200.approaches import Collocation
from ozone_b200.methods import RK4, Trapezoidal
from ozone_b200.optimize import solve_collocation, CollocationNLP

pytestmark = pytest.mark.gpu


def var(value):
    return torch.tensor(np.asarray(value, dtype=np.float64), device='cuda')


@pytest.mark.parametrize('method', [RK4(), Trapezoidal()], ids=['RK4', 'Trapezoidal'])
def test_1_simple_collocation(method):
    p = ozone.ODEProblem(method, Collocation())
    p.add_state('y', initial_condition=var([0.5]), store_history=True)
    p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=var(np.ones(29) * 0.01)))
    p.set_function(ozone.rhs.affine, -1.0)
    before = p.solve()
    assert float(before.states['y'].sum()) == pytest.approx(0.5 * 30)           # initial guess = broadcast IC
    out = solve_collocation(p, ftol=1e-9)
    assert out.optimizer_result.success
    assert float(out.states['y'].sum()) == pytest.approx(13.0239779, rel=1e-5)  # test_1_simple/utils.py:51-52
    nlp = CollocationNLP(p)
    assert np.abs(nlp.constraints(nlp.x0())).max() < 1e-8


def test_collocation_jacobian_matches_finite_differences():
    p = ozone.ODEProblem(RK4(), Collocation())
    p.add_state('y', var([2.0]))
    p.add_state('x', var([2.0]))
    p.add_dynamic_parameter('a', var(1.0 + np.arange(6) / 30.0))
    p.set_timespan(ozone.timespans.StepVector(0.0, var(np.full(5, 0.1))))
    p.set_function(ozone.rhs.lotka_volterra, 0.5)
    p.solve()
    nlp = CollocationNLP(p)
    rng = np.random.default_rng(0)
    x = nlp.x0() * (1.0 + 0.05 * rng.standard_normal(nlp.m))
    J = nlp.jacobian(x)
    assert J.shape == (nlp.m, nlp.m)
    dx = rng.standard_normal(nlp.m)
    e = 1e-6
    fd = (nlp.constraints(x + e * dx) - nlp.constraints(x - e * dx)) / (2 * e)
    assert np.allclose(J @ dx, fd, rtol=1e-6, atol=1e-8)


@pytest.mark.parametrize('method', [RK4()], ids=['RK4'])
def test_2_general_collocation(method):
    nt = 31
    t = np.arange(nt)
    p = ozone.ODEProblem(method, Collocation())
    p.add_state('y', initial_condition=var([2.0]))
    p.add_state('x', initial_condition=var([2.0]))
    p.add_state('z', initial_condition=var([[2.0, 1.0], [-1.0, -3.0]]), store_history=True)
    p.add_dynamic_parameter('a', var(1.0 + t / nt / 5.0))
    p.add_dynamic_parameter('b', var(0.5 + t / nt / 5.0))
    p.add_dynamic_parameter('g', var(2.0 + t / nt / 5.0))
    p.add_dynamic_parameter('e', var(np.array([[0.3, 0.2], [0.1, -2.6]])[None] + (t / nt / 5.0)[:, None, None]))
    p.add_dynamic_parameter('coefficients', var(np.ones(nt) / nt))
    p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=var(np.ones(nt - 1) * 0.001)))
    p.set_function(ozone.rhs.general3, var([0.5]))
    p.solve()
    out = solve_collocation(p, ftol=1e-9)
    pox, poz, poy = out.profile_outputs['x'], out.profile_outputs['z'], out.profile_outputs['y']
    total = pox[-1] + poz[-1] + out.field_outputs['y'][0] + out.states['z'][-1, 0, 1].reshape(1) / 2.0
    total2 = torch.linalg.norm(poy[-1] + poy[0])
    assert float(total) == pytest.approx(2.80611596, rel=1e-4)                   # test_2_general/utils.py:161-162
    assert float(total2) == pytest.approx(22.27526371713729, rel=1e-4)
    assert 'y' not in out.states and tuple(out.states['z'].shape) == (nt, 2, 2)
