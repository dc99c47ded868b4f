"""``ODEProblem``: the object a user builds a solve with.

Call-compatible with the reference's class of the same name (constructor, ``add_state``,
``add_dynamic_parameter``, ``set_timespan``, ``set_function``, ``solve``; behaviour documented in
``ozone/odeproblem.py:13-136`` and exercised by ``ozone/tests/test_api_odeproblem.py``): the same argument
meanings, the same exception types under the same conditions, one ``solve()`` per object.  What differs is
underneath: values are ``torch.Tensor`` and ``solve()`` launches CUDA kernels instead of recording a graph.
"""
from __future__ import annotations

from typing import Callable

import numpy as np

from . import integrators as _integrators
from . import utils as _u
from .approaches import _Approach
from .collections import DynamicParameters, IntegratorOutputs, States
from .func_wrapper import FuncWrapper
from .glms import GLMMethod, get_integration_method
from .timespan import TimeSpan

# approach marker -> integrator (the reference binds the class on the marker, ozone/approaches.py:9-30)
_INTEGRATOR_OF = {
    'time_marching': _integrators.TimeMarching,
    'time_marching_checkpoints': _integrators.TimeMarchingUniform,
    'picard': _integrators.PicardIteration,
    'collocation': _integrators.Collocation,
}


def _reject(kind, what):
    raise kind(what)


class ODEProblem(object):
    """``ODEProblem(method, approach, **engine_options)``

    engine_options (all optional, none exist in the reference): ``newton_tolerance`` (1e-7 in fp64, the value the
    reference hard-codes at time_marching.py:220; 1e-5 for fp32 problems, whose residual cannot get below ~1e-7), ``newton_max_iter`` (100), ``verbose`` (checkpoint banner),
    ``batched`` (Python callbacks only: the trailing axis of every state is the trajectory axis).
    """

    def __init__(self, method, approach, **engine_options) -> None:
        _u.check_if_method(method, 'method')                       # TypeError for anything but a methods.* instance
        if not isinstance(approach, _Approach):
            _reject(TypeError, f"argument 'approach' must be an ozone approach instance, got {_u.get_type_string(approach)}")
        self.method: GLMMethod = get_integration_method(method.name)
        self.approach = approach
        self.options = dict(engine_options)
        self.states, self.d_params = States(), DynamicParameters()
        self.f: FuncWrapper = None
        self.time_vector = self.step_vector = self.num_times = None
        self.solved = False

    # ---- problem definition ---------------------------------------------------------------------------
    def add_state(self, name: str, initial_condition, store_history: bool = False, store_final: bool = False,
                  initial_guess: np.ndarray = None, scaler=None) -> None:
        """Register a state.  ``store_history`` / ``store_final`` decide what ``solve()`` returns for it;
        ``initial_guess`` (num_times-1, *shape) seeds Picard / Collocation; ``scaler`` scales the collocation
        design variables."""
        _u.check_if_string(name, 'name')
        ic = _u.variablize(initial_condition, 'initial_condition')
        for flag, label in ((store_history, 'store_history'), (store_final, 'store_final')):
            _u.check_if_bool(flag, label)
        if not (initial_guess is None or isinstance(initial_guess, np.ndarray)):
            _reject(TypeError, f"argument 'initial_guess' must be a numpy array, got {_u.get_type_string(initial_guess)}")
        self.states.add(name, ic, store_history, store_final, initial_guess, scaler)      # ValueError on a duplicate

    def add_dynamic_parameter(self, name: str, parameter) -> None:
        """Register a parameter given at every time point (first axis = num_times, checked in ``solve``)."""
        _u.check_if_string(name, 'name')
        self.d_params.add(name, _u.variablize(parameter, 'parameter'))                    # ValueError on a duplicate

    def set_timespan(self, timespan: TimeSpan) -> None:
        if not isinstance(timespan, TimeSpan):
            _reject(TypeError, f"argument 'timespan' must be an ozone timespan, got {_u.get_type_string(timespan)}")
        if self.num_times is not None:
            _reject(ValueError, 'the timespan of this problem is already set')
        self.time_vector, self.step_vector, self.num_times = timespan.finalize()

    def set_function(self, f: Callable, *args, **kwargs) -> None:
        """``f(ozone_vars, *args, **kwargs)``: a registered device right-hand side (``ozone_b200.rhs.*``) or a
        Python callback, which is traced and compiled on ``solve()``."""
        if not callable(f):
            _reject(TypeError, f"argument 'f' must be callable, got {_u.get_type_string(f)}")
        self.f = FuncWrapper(f, args, kwargs)

    # ---- solve ------------------------------------------------------------------------------------------
    def solve(self) -> IntegratorOutputs:
        problems = (
            (self.solved, 'this problem has been solved already; build a new ODEProblem to solve again'),
            (self.f is None, "no ODE function: call set_function(f) before solve()"),
            (self.num_times is None, 'no time span: call set_timespan(...) before solve()'),
        )
        for bad, message in problems:
            if bad:
                _reject(ValueError, message)
        self.states.validate(self.num_times)                       # ValueError: no states / wrong initial-guess shape
        self.d_params.validate(self.num_times)                     # ValueError: first axis != num_times
        self.solved = True                                         # as in the reference: a failed solve is not retried
        cls = _INTEGRATOR_OF[self.approach.kind]
        self.integrator = cls(method=self.method, approach=self.approach, ode_func=self.f, states=self.states,
                              dynamic_parameters=self.d_params, num_times=self.num_times,
                              step_vector=self.step_vector, options=self.options)
        self.integrator.integrate()
        return self.integrator.finalize_outputs()
