"""The RHS plugin boundary (``ozone/func_wrapper.py``): ``ODEVars`` is what the callback sees,
``FuncWrapper.evaluate`` calls it and validates what it set."""
from typing import Callable

import torch


class ODEVars(object):
    def __init__(self, states, dynamic_parameters, num_nodes) -> None:
        self.states = states
        self.dynamic_parameters = dynamic_parameters
        self.num_nodes = num_nodes
        for state_name, state_input in self.states.items():
            if state_input.shape[0] != num_nodes:
                raise ValueError(f'INTERNAL ERROR: Expected shape ({num_nodes}, ...) for state \'{state_name}\' but got {tuple(state_input.shape)}')
        for name, dp in self.dynamic_parameters.items():
            if dp.shape[0] != num_nodes:
                raise ValueError(f'INTERNAL ERROR: Expected shape ({num_nodes}, ...) for dynamic parameter \'{name}\' but got {tuple(dp.shape)}')
        self.d_states = {}
        self.profile_outputs = {}
        self.field_outputs = {}

    def post_process(self):
        """ozone/func_wrapper.py:28-53: KeyError for missing / extra derivative keys, ValueError for
        wrong shapes, same-size derivatives are reshaped, 1-D outputs become (n, 1)."""
        for state_name, state_input in self.states.items():
            if state_name not in self.d_states:
                raise KeyError(f'No state derivative specified for state \'{state_name}\'.')
            ds = self.d_states[state_name]
            if ds.shape != state_input.shape:
                if ds.numel() == state_input.numel():
                    self.d_states[state_name] = ds.reshape(state_input.shape)
                else:
                    raise ValueError(f'Expected shape {tuple(state_input.shape)} for state derivative \'{state_name}\' but got {tuple(ds.shape)}')
        for d_state_name in self.d_states:
            if d_state_name not in self.states:
                raise KeyError(f'User set state derivative \'{d_state_name}\' but state \'{d_state_name}\' does not exist. Given states: {list(self.states.keys())}')
        for kind, outs in (('profile', self.profile_outputs), ('field', self.field_outputs)):
            for output_name, out in list(outs.items()):
                if out.dim() == 0 or out.shape[0] != self.num_nodes:
                    raise ValueError(f'{kind} output {output_name} expects shape ({self.num_nodes}, ...) but got {tuple(out.shape)}')
                if out.dim() < 2:
                    outs[output_name] = out.reshape(self.num_nodes, 1)

    def __str__(self):
        groups = [('States:', self.states), ('Dynamic Parameters:', self.dynamic_parameters),
                  ('F(states):', self.d_states), ('Profile Outputs:', self.profile_outputs),
                  ('Field Outputs:', self.field_outputs)]
        string = '\nODE Function inputs/outputs:\n'
        for title, vars_ in groups:
            string += f'{title}\n'
            for key, value in vars_.items():
                string += f'\t{key}: {tuple(value.shape)}\n'
        return string


class FuncWrapper(object):
    def __init__(self, f: Callable, args: tuple, kwargs: dict) -> None:
        self.f = f
        self.args, self.kwargs = args, kwargs
        self.profile_shapes = {}
        self.field_shapes = {}
        self.ran_once = False

    @property
    def is_device(self):
        from .rhs import DeviceRHS
        return isinstance(self.f, DeviceRHS)

    def evaluate(self, states: dict, parameters: dict, num_nodes: int) -> ODEVars:
        """ozone/func_wrapper.py:90-107."""
        func_vars = ODEVars(states, parameters, num_nodes)
        self.f(func_vars, *self.args, **self.kwargs)
        func_vars.post_process()
        if not self.ran_once:
            for name, v in func_vars.profile_outputs.items():
                self.profile_shapes[name] = tuple(v.shape[1:])
            for name, v in func_vars.field_outputs.items():
                self.field_shapes[name] = tuple(v.shape[1:])
        self.ran_once = True
        return func_vars
