"""``ODEProblem``: the user-facing problem object, API-compatible with ``ozone/odeproblem.py:13-136``."""
from __future__ import annotations

from typing import Callable

import numpy as np

from . import integrators as _integrators
from .approaches import _Approach
from .collections import DynamicParameters, IntegratorOutputs, States
from .func_wrapper import FuncWrapper
from .glms import GLMMethod, get_integration_method
from .timespan import TimeSpan
from .utils import check_if_bool, check_if_method, check_if_string, get_type_string, variablize

_INTEGRATORS = {
    'time_marching': _integrators.TimeMarching,
    'time_marching_checkpoints': _integrators.TimeMarchingUniform,
    'picard': _integrators.PicardIteration,
    'collocation': _integrators.Collocation,
}


class ODEProblem(object):

    def __init__(self, method, approach, **options) -> None:
        """``ODEProblem(method, approach)`` as in the reference.  Optional keyword options of the
        device engine: ``newton_tolerance`` (default 1e-7, time_marching.py:220), ``newton_max_iter``,
        ``verbose``, ``batched`` (Python callbacks only: the trailing axis of every state is the
        trajectory axis; registered ``rhs.*`` handles infer it from their declared shapes)."""
        check_if_method(method, 'method')
        self.method: GLMMethod = get_integration_method(method.name)
        if not isinstance(approach, _Approach):
            raise TypeError(f'expected type \'_Approach\' for argument \'approach\' but got {get_type_string(approach)}')
        self.approach = approach
        self.options = options
        self.states = States()
        self.d_params = DynamicParameters()
        self.solved = False
        self.time_vector = None
        self.step_vector = None
        self.num_times = None
        self.f: FuncWrapper = None

    def add_state(self, name: str, initial_condition, store_history: bool = False, store_final: bool = False,
                  initial_guess: np.ndarray = None, scaler=None) -> None:
        check_if_string(name, 'name')
        initial_condition = variablize(initial_condition, 'initial_condition')
        check_if_bool(store_history, 'store_history')
        check_if_bool(store_final, 'store_final')
        if initial_guess is not None and not isinstance(initial_guess, np.ndarray):
            raise TypeError(f'expected type \'np.ndarray\' for argument \'initial_guess\' but got {get_type_string(initial_guess)}')
        self.states.add(name, initial_condition, store_history, store_final, initial_guess, scaler)

    def add_dynamic_parameter(self, name: str, parameter) -> None:
        check_if_string(name, 'name')
        parameter = variablize(parameter, 'parameter')
        self.d_params.add(name, parameter)

    def set_timespan(self, timespan: TimeSpan) -> None:
        if not isinstance(timespan, TimeSpan):
            raise TypeError(f'expected type \'TimeSpan\' for argument \'timespan\' but got {get_type_string(timespan)}')
        if self.num_times is not None:
            raise ValueError('timespan has already been set.')
        self.time_vector, self.step_vector, self.num_times = timespan.finalize()

    def set_function(self, f: Callable, *args, **kwargs) -> None:
        if not callable(f):
            raise TypeError(f'expected type \'callable\' for argument \'f\' but got {get_type_string(f)}')
        self.f = FuncWrapper(f, args, kwargs)

    def solve(self) -> IntegratorOutputs:
        if self.solved:
            raise ValueError('solve has already been called.')
        if self.f is None:
            raise ValueError('The function \'f\' must be set using the \'set_function\' method before solving the ODE.')
        if self.num_times is None:
            raise ValueError('The integration time span must be set using the \'set_timespan\' method before solving the ODE.')
        self.states.validate(self.num_times)
        self.d_params.validate(self.num_times)
        self.solved = True
        self.integrator = _INTEGRATORS[self.approach.kind](
            method=self.method, approach=self.approach, ode_func=self.f, states=self.states,
            dynamic_parameters=self.d_params, num_times=self.num_times, step_vector=self.step_vector,
            options=self.options)
        self.integrator.integrate()
        return self.integrator.finalize_outputs()
