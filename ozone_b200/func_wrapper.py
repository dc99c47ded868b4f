"""The right-hand side plugin boundary.

``ODEVars`` is the bag of named arrays a callback receives and fills (same attribute names as the reference's
``ozone/func_wrapper.py:5-26``: ``states``, ``dynamic_parameters``, ``num_nodes`` in; ``d_states``,
``profile_outputs``, ``field_outputs`` out).  ``FuncWrapper`` holds the callback with its extra arguments and
checks what it produced against the contract the reference enforces (``func_wrapper.py:28-53``): a derivative
for every state and for nothing else (KeyError), shapes that match or at least sizes that match (ValueError
otherwise; equal sizes are reshaped), outputs whose leading axis is ``num_nodes`` (1-D outputs become columns).
"""
from typing import Callable


def _size(x):
    return x.numel() if hasattr(x, 'numel') else x.size


class ODEVars(object):
    def __init__(self, states, dynamic_parameters, num_nodes) -> None:
        self.states, self.dynamic_parameters, self.num_nodes = states, dynamic_parameters, num_nodes
        for kind, group in (('state', states), ('dynamic parameter', dynamic_parameters)):
            for name, arr in group.items():
                if arr.shape[0] != num_nodes:
                    raise ValueError(f"internal: {kind} '{name}' was handed over with shape {tuple(arr.shape)}, "
                                     f"its leading axis should be num_nodes = {num_nodes}")
        self.d_states, self.profile_outputs, self.field_outputs = {}, {}, {}

    # ---- checks run after the callback returned
    def _check_derivatives(self):
        missing = [n for n in self.states if n not in self.d_states]
        if missing:
            raise KeyError(f"the ODE function set no derivative for state '{missing[0]}'")
        unknown = [n for n in self.d_states if n not in self.states]
        if unknown:
            raise KeyError(f"the ODE function set a derivative for '{unknown[0]}', which is not a state "
                           f"(states: {list(self.states)})")
        for name, y in self.states.items():
            dy = self.d_states[name]
            if tuple(dy.shape) == tuple(y.shape):
                continue
            if _size(dy) != _size(y):
                raise ValueError(f"derivative of state '{name}' has shape {tuple(dy.shape)}; expected {tuple(y.shape)}")
            self.d_states[name] = dy.reshape(tuple(y.shape))       # same size, other shape: accepted and reshaped

    def _check_outputs(self, label, outs):
        for name, val in list(outs.items()):
            shape = tuple(val.shape)
            if len(shape) == 0 or shape[0] != self.num_nodes:
                raise ValueError(f"{label} output '{name}' has shape {shape}; its leading axis must be num_nodes = {self.num_nodes}")
            if len(shape) == 1:
                outs[name] = val.reshape(self.num_nodes, 1)

    def post_process(self):
        self._check_derivatives()
        self._check_outputs('profile', self.profile_outputs)
        self._check_outputs('field', self.field_outputs)

    def __str__(self):
        sections = (('States', self.states), ('Dynamic parameters', self.dynamic_parameters),
                    ('State derivatives', self.d_states), ('Profile outputs', self.profile_outputs),
                    ('Field outputs', self.field_outputs))
        lines = ['', 'ODE function variables:']
        for title, group in sections:
            lines.append(f'  {title}:')
            lines += [f'    {k}: {tuple(v.shape)}' for k, v in group.items()]
        return '\n'.join(lines) + '\n'


class FuncWrapper(object):
    """The callback plus the ``*args, **kwargs`` given to ``set_function``; remembers the output shapes it saw."""

    def __init__(self, f: Callable, args: tuple, kwargs: dict) -> None:
        self.f, self.args, self.kwargs = f, args, kwargs
        self.profile_shapes, self.field_shapes = {}, {}
        self.ran_once = False

    @property
    def is_device(self):
        from .rhs import DeviceRHS
        return isinstance(self.f, DeviceRHS)

    def evaluate(self, states: dict, parameters: dict, num_nodes: int) -> ODEVars:
        """Call ``f`` on ``num_nodes`` rows and validate the result (``ozone/func_wrapper.py:90-107``)."""
        ov = ODEVars(states, parameters, num_nodes)
        self.f(ov, *self.args, **self.kwargs)                      # a wrong argument list is a plain TypeError
        ov.post_process()
        if not self.ran_once:                                      # outputs are discovered on the first call
            self.profile_shapes = {k: tuple(v.shape[1:]) for k, v in ov.profile_outputs.items()}
            self.field_shapes = {k: tuple(v.shape[1:]) for k, v in ov.field_outputs.items()}
            self.ran_once = True
        return ov
