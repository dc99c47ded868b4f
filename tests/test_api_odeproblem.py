"""API / error-path tests, mirroring ozone/tests/test_api_odeproblem.py case by case
(line numbers in comments refer to that file).  They need no GPU: every error below is raised on
the host before any kernel is launched."""
import numpy as np
import pytest
import torch

import ozone_b200 as ozone


def _t(v):
    return torch.tensor(v, dtype=torch.float64)


def test_ode_problem_args():
    with pytest.raises(TypeError):                                   # :10-11
        ozone.ODEProblem(ozone.methods.RK4(), 1)
    with pytest.raises(TypeError):                                   # :13-14
        ozone.ODEProblem(1, 'picard_iteration')
    with pytest.raises(TypeError):                                   # method class instead of instance
        ozone.ODEProblem(ozone.methods.RK4, ozone.approaches.TimeMarching())

    picard_iteration = ozone.approaches.PicardIteration()
    ode_problem = ozone.ODEProblem(ozone.methods.RK4(), picard_iteration)
    x0 = _t([2.0])
    with pytest.raises(TypeError):                                   # :21-23 name not str
        ode_problem.add_state(1, x0)
    with pytest.raises(TypeError):                                   # :24-26 IC not variable-like
        ode_problem.add_state('x', 'a')
    with pytest.raises(TypeError):                                   # :27-29 store_history not bool
        ode_problem.add_state('x', x0, store_history=1)
    with pytest.raises(ValueError):                                  # :30-33 duplicate state
        ode_problem.add_state('x', x0, store_history=True)
        ode_problem.add_state('x', x0 * 1.0)
    with pytest.raises(TypeError):
        ode_problem.add_state('w', x0, initial_guess=[1.0])          # odeproblem.py:56-58

    a = _t([2.0])
    with pytest.raises(TypeError):                                   # :36-38
        ode_problem.add_dynamic_parameter(1, a)
    with pytest.raises(TypeError):                                   # :39-41
        ode_problem.add_dynamic_parameter('a', 'a')
    with pytest.raises(ValueError):                                  # :42-45 duplicate parameter
        ode_problem.add_dynamic_parameter('a', a)
        ode_problem.add_dynamic_parameter('a', a * 1.0)

    with pytest.raises(TypeError):                                   # :48-49
        ode_problem.set_timespan('a')
    with pytest.raises(ValueError):                                  # :50-52 timespan set twice
        ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=1.0))
        ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=1.0))
    with pytest.raises(TypeError):                                   # odeproblem.py:99-100 (untested in the reference)
        ode_problem.set_function(3)

    # dynamic parameter first axis != num_times                      # :55-62
    ode_problem = ozone.ODEProblem(ozone.methods.RK4(), picard_iteration)
    ode_problem.add_state('x', x0)
    ode_problem.add_dynamic_parameter('a', np.array([1.0, 2.0]))
    ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=[0.0, 1.0]))
    ode_problem.set_function(lambda x: None)
    with pytest.raises(ValueError):
        ode_problem.solve()

    # no ODE function                                                # :64-70
    ode_problem = ozone.ODEProblem(ozone.methods.RK4(), picard_iteration)
    ode_problem.add_state('x', x0)
    ode_problem.add_dynamic_parameter('a', np.array([1.0, 2.0]))
    ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=[0.0, 1.0]))
    with pytest.raises(ValueError):
        ode_problem.solve()

    # no timespan                                                    # :72-78
    ode_problem = ozone.ODEProblem(ozone.methods.RK4(), picard_iteration)
    ode_problem.add_state('x', x0)
    ode_problem.add_dynamic_parameter('a', np.array([1.0, 2.0]))
    ode_problem.set_function(lambda x: None)
    with pytest.raises(ValueError):
        ode_problem.solve()

    # no states                                                      # :80-86
    ode_problem = ozone.ODEProblem(ozone.methods.RK4(), picard_iteration)
    ode_problem.add_dynamic_parameter('a', np.array([1.0, 2.0]))
    ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=[0.0, 1.0]))
    ode_problem.set_function(lambda x: None)
    with pytest.raises(ValueError):
        ode_problem.solve()

    # solve twice                                                    # :88-100
    ode_problem = ozone.ODEProblem(ozone.methods.RK4(), ozone.approaches.TimeMarching())
    ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=[0.0, 1.0]))
    ode_problem.set_function(ozone.rhs.affine, 1.0)
    ode_problem.add_state('y', x0)
    try:
        ode_problem.solve()
    except RuntimeError:            # no CUDA device in this process: the first solve is refused (no CPU fallback)
        assert not torch.cuda.is_available()
    with pytest.raises(ValueError):
        ode_problem.solve()


def _problem_with(f, *args, **kwargs):
    num_times = 10
    ode_problem = ozone.ODEProblem(ozone.methods.RK4(), ozone.approaches.PicardIteration())
    ode_problem.add_state('x', torch.ones((1, 2), dtype=torch.float64))
    ode_problem.add_dynamic_parameter('a_dynamic', torch.ones((num_times, 3), dtype=torch.float64))
    ode_problem.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=0.1 * np.ones((num_times - 1,))))
    ode_problem.set_function(f, *args, **kwargs)
    return ode_problem


def _expect_traced(problem):
    """A well-formed Python callback is traced into a device functor; what happens next depends on the box:
    with a GPU it solves, without one the solve is refused (no CPU fallback) -- never NotImplementedError."""
    try:
        problem.solve()
    except RuntimeError as e:
        assert not torch.cuda.is_available() and 'CUDA' in str(e)


def test_ode_function_api_python_callbacks_are_validated_like_the_reference():
    a_static = torch.ones((1,), dtype=torch.float64)

    def rate(v):
        return -v.states['x'] * v.dynamic_parameters['a_dynamic'].sum(axis=1).reshape(-1, 1, 1) * a_static

    # a well-formed callback is accepted (traced to device code)                                   # :113-124
    def f(v: ozone.ODEVars):
        v.d_states['x'] = rate(v)
        v.field_outputs['x'] = v.states['x']
        v.profile_outputs['x'] = v.states['x']
    _expect_traced(_problem_with(f))

    def f(v):                                                        # :127-140 extra key
        v.d_states['x'] = rate(v)
        v.d_states['x1'] = rate(v)
    with pytest.raises(KeyError):
        _problem_with(f).solve()

    def f(v):                                                        # :142-154 key not found
        v.d_states['x1'] = rate(v)
    with pytest.raises(KeyError):
        _problem_with(f).solve()

    def f(v):                                                        # :156-169 state shape wrong
        v.d_states['x'] = rate(v)[0, 0]
    with pytest.raises(ValueError):
        _problem_with(f).solve()

    def f(v):                                                        # same size, different shape: reshaped (func_wrapper.py:32-34)
        v.d_states['x'] = rate(v).reshape(-1, 2)
    _expect_traced(_problem_with(f))

    def f(v):                                                        # :171-184 profile output shape wrong
        v.d_states['x'] = rate(v)
        v.profile_outputs['x'] = v.states['x'][0, :, :].reshape(2)
    with pytest.raises(ValueError):
        _problem_with(f).solve()

    def f(v):                                                        # :186-199 field output shape wrong
        v.d_states['x'] = rate(v)
        v.field_outputs['x'] = v.states['x'][0, :, :].reshape(2)
    with pytest.raises(ValueError):
        _problem_with(f).solve()

    def f(v, d):                                                     # :201-213 args wrong
        v.d_states['x'] = rate(v) * d
    with pytest.raises(TypeError):
        _problem_with(f).solve()

    def f(v, d):                                                     # :215-227 kwargs wrong
        v.d_states['x'] = rate(v)
    with pytest.raises(TypeError):
        _problem_with(f, d1=1).solve()


def _lv_problem(**kw):
    nt = 5
    p = ozone.ODEProblem(ozone.methods.RK4(), ozone.approaches.TimeMarching())
    for n in kw.get('states', ('y', 'x')):
        p.add_state(n, _t([2.0]))
    for n in kw.get('params', ('a',)):
        p.add_dynamic_parameter(n, np.ones(nt))
    p.set_timespan(ozone.timespans.StepVector(start=0.0, step_vector=np.full(nt - 1, 0.1)))
    p.set_function(ozone.rhs.lotka_volterra, *kw.get('args', (0.5,)), **kw.get('kwargs', {}))
    return p


def test_device_rhs_contract_errors_match_the_callback_contract():
    with pytest.raises(KeyError):           # the function reads a state that was not added
        _lv_problem(states=('y',)).solve()
    with pytest.raises(KeyError):           # a state without a derivative (func_wrapper.py:30-31)
        _lv_problem(states=('y', 'x', 'w')).solve()
    with pytest.raises(KeyError):           # missing dynamic parameter
        _lv_problem(params=()).solve()
    with pytest.raises(TypeError):          # missing positional argument
        _lv_problem(args=()).solve()
    with pytest.raises(TypeError):          # too many
        _lv_problem(args=(0.5, 1.0)).solve()
    with pytest.raises(TypeError):          # unknown keyword
        _lv_problem(args=(), kwargs={'d1': 1}).solve()
    p = ozone.ODEProblem(ozone.methods.RK4(), ozone.approaches.TimeMarching())
    p.add_state('y', torch.ones((3, 3), dtype=torch.float64))       # wrong point shape
    p.add_state('x', _t([2.0]))
    p.add_dynamic_parameter('a', np.ones(3))
    p.set_timespan(ozone.timespans.StepVector(0.0, np.full(2, 0.1)))
    p.set_function(ozone.rhs.lotka_volterra, 0.5)
    with pytest.raises(ValueError):
        p.solve()
    p = ozone.ODEProblem(ozone.methods.RK4(), ozone.approaches.TimeMarching())
    p.add_state('y', torch.ones(4, dtype=torch.float64))            # batch of 4 ...
    p.add_state('x', torch.ones(5, dtype=torch.float64))            # ... and of 5
    p.add_dynamic_parameter('a', np.ones(3))
    p.set_timespan(ozone.timespans.StepVector(0.0, np.full(2, 0.1)))
    p.set_function(ozone.rhs.lotka_volterra, 0.5)
    with pytest.raises(ValueError):
        p.solve()


def test_checkpoints_divisibility_and_initial_guess_shape():
    p = ozone.ODEProblem(ozone.methods.RK4(), ozone.approaches.TimeMarchingCheckpoints(num_checkpoints=5))
    p.add_state('y', _t([0.5]))
    p.set_timespan(ozone.timespans.StepVector(0.0, np.full(29, 0.01)))   # test_1_simple: 29 steps, nc = 5
    p.set_function(ozone.rhs.affine, -1.0)
    with pytest.raises(ValueError):                                      # time_marching_uniform.py:25-26
        p.solve()
    p = ozone.ODEProblem(ozone.methods.RK4(), ozone.approaches.Collocation())
    p.add_state('y', _t([0.5]), initial_guess=np.ones((7, 1)))           # collections.py:92-95
    p.set_timespan(ozone.timespans.StepVector(0.0, np.full(29, 0.01)))
    p.set_function(ozone.rhs.affine, -1.0)
    with pytest.raises(ValueError):
        p.solve()


def test_timespans():
    ts = ozone.timespans.StepVector(start=0.0, step_vector=np.full(9, 0.1))
    tv, h, nt = ts.finalize()
    assert tv is None and nt == 10 and tuple(h.shape) == (9,)            # timespan.py:25-33
    with pytest.raises(ValueError):
        ozone.timespans.StepVector(start=[0.0, 1.0], step_vector=np.full(9, 0.1))      # start not scalar
    with pytest.raises(ValueError):
        ozone.timespans.StepVector(start=0.0, step_vector=np.ones((2, 2, 2)))          # not a vector
    with pytest.raises(TypeError):
        ozone.timespans.StartEnd(0.0, 1.0, 2.5)
    tv, h, nt = ozone.timespans.StartEnd(0.0, 1.0, 4).finalize()
    assert nt == 5 and torch.allclose(h.cpu(), torch.full((4,), 0.25, dtype=h.dtype))
    assert ozone.timespan is ozone.timespans                              # both spellings


def test_methods_and_approaches_surface():
    names = ['ForwardEuler', 'BackwardEuler', 'ExplicitMidpoint', 'ImplicitMidpoint', 'KuttaThirdOrder', 'RK4', 'RK6',
             'RalstonsMethod', 'HeunsMethod', 'GaussLegendre2', 'GaussLegendre4', 'GaussLegendre6', 'Lobatto2',
             'Lobatto4', 'RadauI3', 'RadauI5', 'RadauII3', 'RadauII5', 'Trapezoidal', 'AB1', 'AM1', 'BDF1']
    for n in names:                                                       # ozone/methods.py
        m = getattr(ozone.methods, n)()
        assert isinstance(m, ozone.methods._Method) and m.name == n
    ap = ozone.approaches                                                 # ozone/approaches.py:12,17,25,30
    assert ap.TimeMarching().name == 'TimeMarching'
    assert ap.TimeMarchingCheckpoints().name == 'TimeMarching (Checkpointing)'
    assert ap.TimeMarchingCheckpoints(num_checkpoints=7).num_checkpoints == 7
    assert ap.PicardIteration().name == 'Picard Iteration' and ap.Collocation().name == 'Collocation'


def test_python_callback_tracing_rules():
    from ozone_b200 import trace, ops
    sp = torch.linspace(0.1, 1.1, 3, dtype=torch.float64)

    def f(v):                                   # test_3_large/utils.py:23-30 written for one trajectory
        y, dp = v.states['y'], v.dynamic_parameters['dynamic_param']
        v.d_states['y'] = -(y * (dp + sp[0])) ** 2
    t = trace.TracedRHS(f, (), {}, {'y': ()}, {'dynamic_param': ()})
    assert (t.D, t.P, t.S, t.DQ) == (1, 1, 1, 0) and 'struct RhsTraced' in t.cuda_source()

    def vdp(v):                                 # van_der_pol/main_model.py:6-14: J is a pure quadrature
        x0, x1, u = v.states['x0'], v.states['x1'], v.dynamic_parameters['u']
        v.d_states['x0'] = ((1.0 + (-1.0) * (x1 ** 2.0)) * x0 + (-1 * x1) + u).reshape(-1, 1)
        v.d_states['x1'] = (x0 * 1.0).reshape(-1, 1)
        v.d_states['J'] = (x0 ** 2.0 + x1 ** 2.0 + u ** 2.0).reshape(-1, 1)
    t = trace.TracedRHS(vdp, (), {}, {'x0': (), 'x1': (), 'J': ()}, {'u': ()})
    assert (t.D, t.P, t.S, t.DQ) == (3, 1, 0, 1)

    def couples_rows(v):                        # not node-wise: every row reads row 0
        v.d_states['y'] = v.states['y'][0:1].expand(v.states['y'].shape)
    with pytest.raises(ValueError):
        trace.TracedRHS(couples_rows, (), {}, {'y': ()}, {})

    def branches(v):                            # data-dependent Python control flow cannot be compiled
        y = v.states['y']
        v.d_states['y'] = y if y[0, 0] > 0 else -y
    with pytest.raises(TypeError):
        trace.TracedRHS(branches, (), {}, {'y': ()}, {})

    def big(v):
        v.d_states['y'] = -v.states['y']
    with pytest.raises(ValueError):             # 500 components per trajectory do not fit one thread: batch them
        trace.TracedRHS(big, (), {}, {'y': (500,)}, {})

    def uses_ops(v):
        y = v.states['y']
        v.d_states['y'] = ops.exp(-y) * ops.sin(y) + np.cos(y) / ops.sqrt(1.0 + y * y) - abs(y) ** 1.5
    t = trace.TracedRHS(uses_ops, (), {}, {'y': (2,)}, {})
    src = t.cuda_source()
    assert 'exp(' in src and 'sin(' in src and 'cos(' in src and 'sqrt(' in src and 'pow(' in src
