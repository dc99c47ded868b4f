"""Registered device right-hand sides: what ``set_function`` accepts.

The reference hands a Python callback ``f(ozone_vars, *args, **kwargs)`` to ``set_function``
(``ozone/odeproblem.py:93-101``) and csdl records/differentiates it.  A Python callback cannot run
inside a CUDA kernel, so each named problem of the reference is restated as a device functor
(``csrc/oz_rhs.cuh``) and exposed here as a callable handle with the same calling convention:

    ode_problem.set_function(ozone_b200.rhs.lotka_volterra, c_static)

States / dynamic parameters are matched BY NAME with what was added to the problem, exactly as the
reference callback would index ``ozone_vars.states[name]`` (a missing name is a KeyError).  Extra
positional / keyword arguments are the callback's static ("closure-captured") parameters; they
are differentiable inputs.  Any other (node-wise) Python callback goes through ``ozone_b200/trace.py``: it is run once
on symbolic inputs, differentiated, emitted as a device functor of the same contract and compiled into a plugin
(same error behaviour as ``ozone/func_wrapper.py`` for malformed callbacks); the handles here skip that step.
"""
from __future__ import annotations

import inspect
import math

import numpy as np

from . import _native as nat


class DeviceRHS:
    """Handle of one device functor.  Lists are (name, point_shape) in PACKED order."""
    kernel = ''
    kind = 'batched'                       # 'batched' (one ODE per trajectory) | 'grid'
    states = ()
    dynamic_parameters = ()
    static_parameters = ()                 # bound from set_function(f, *args, **kwargs)
    static_defaults = {}
    profile_outputs = ()
    field_outputs = ()
    config_parameters = ()                 # non-differentiable keyword constants (grid RHS)

    def __init__(self, rename=None):
        self._rename = dict(rename or {})

    def _n(self, name):
        return self._rename.get(name, name)

    def names(self, which):
        return [(self._n(n), None if s is None else tuple(s)) for n, s in getattr(self, which)]

    @property
    def rhs_id(self):
        return nat.RHS_IDS[self.kernel]

    def bind_static(self, args, kwargs):
        """Python-call semantics for the extra arguments: wrong count / unknown keyword -> TypeError."""
        params = [inspect.Parameter('ozone_vars', inspect.Parameter.POSITIONAL_OR_KEYWORD)]
        for n, _ in list(self.static_parameters) + [(c, ()) for c in self.config_parameters]:
            default = self.static_defaults.get(n, inspect.Parameter.empty)
            params.append(inspect.Parameter(n, inspect.Parameter.POSITIONAL_OR_KEYWORD, default=default))
        bound = inspect.Signature(params).bind(None, *args, **kwargs)     # raises TypeError like a real call
        bound.apply_defaults()
        vals = dict(bound.arguments)
        vals.pop('ozone_vars')
        return vals

    def rhs_cfg(self, static_values, states):
        return ()

    def __call__(self, ozone_vars, *args, **kwargs):
        """Evaluate on the device at ``ozone_vars.num_nodes`` nodes (FuncWrapper.evaluate)."""
        from .integrators import evaluate_device_rhs
        evaluate_device_rhs(self, ozone_vars, self.bind_static(args, kwargs))

    def __repr__(self):
        return f'<device RHS {self.kernel!r}>'


class LotkaVolterra(DeviceRHS):
    """examples/simple_examples/lotka_volterra.py:51-62:  y' = a y - 0.5 y x,  x' = 2 x y - c x;
    profile 'x_plus_y', field 'field_x'.  ``set_function(rhs.lotka_volterra, c)``."""
    kernel = 'lotka_volterra'
    states = (('y', ()), ('x', ()))
    dynamic_parameters = (('a', ()),)
    static_parameters = (('c', ()),)
    profile_outputs = (('x_plus_y', ()),)
    field_outputs = (('field_x', ()),)


class VanDerPol(DeviceRHS):
    """ozone/paper_examples/van_der_pol/main_model.py:6-14: states x0, x1, J; control u."""
    kernel = 'van_der_pol'
    states = (('x0', ()), ('x1', ()), ('J', ()))
    dynamic_parameters = (('u', ()),)


class LunarAscent(DeviceRHS):
    """ozone/paper_examples/ascent_system/main_model.py:6-21."""
    kernel = 'lunar_ascent'
    states = tuple((n, ()) for n in ('rx', 'ry', 'rz', 'Vx', 'Vy', 'Vz', 'm'))
    dynamic_parameters = (('ux', ()), ('uy', ()), ('uz', ()))


class Affine(DeviceRHS):
    """y' = rate*y + offset.  test_1_simple/utils.py:6-7 is ``set_function(rhs.affine, -1.0, 0.0)``;
    the tutorial notebook's ``0.5*y - 3`` is ``set_function(rhs.affine, 0.5, -3.0)``."""
    kernel = 'affine'
    states = (('y', ()),)
    static_parameters = (('rate', ()), ('offset', ()))
    static_defaults = {'offset': 0.0}


class QuadDecay(DeviceRHS):
    """ozone/tests/test_correctness/test_3_large/utils.py:23-30: y' = -(y (dynamic_param + static_param))^2."""
    kernel = 'quad_decay'
    states = (('y', ()),)
    dynamic_parameters = (('dynamic_param', ()),)
    static_parameters = (('static_param', ()),)


class General3(DeviceRHS):
    """ozone/tests/test_correctness/test_2_general/utils.py:43-84 (three states incl. a 2x2 one, five
    dynamic parameters, one static argument, three profile and two field outputs)."""
    kernel = 'general3'
    states = (('y', ()), ('x', ()), ('z', (2, 2)))
    dynamic_parameters = (('a', ()), ('b', ()), ('g', ()), ('e', (2, 2)), ('coefficients', ()))
    static_parameters = (('d_arg', ()),)
    profile_outputs = (('z', ()), ('x', ()), ('y', (2, 2)))
    field_outputs = (('y', ()), ('z', (2, 2)))


class DiffusionReaction(DeviceRHS):
    """ozone/paper_examples/pde_control/pde_control.py:11-81: one trajectory whose state 'x' is the
    grid (nx cells), control 'u', field output 'J_space'.  Keyword constants as in the reference:
    ``set_function(rhs.diffusion_reaction, dx=dx, nx=nx, zero_bc=0, pi_bc=0, beta_t=16., beta_u=2., gamma=2.)``."""
    kernel = 'diffusion_reaction'
    kind = 'grid'
    states = (('x', None),)                    # point shape (nx,) taken from the initial condition
    dynamic_parameters = (('u', ()),)
    field_outputs = (('J_space', ()),)
    config_parameters = ('dx', 'nx', 'zero_bc', 'pi_bc', 'beta_t', 'beta_u', 'gamma')
    static_defaults = {'zero_bc': 0, 'pi_bc': 0, 'beta_t': 16.0, 'beta_u': 2.0, 'gamma': 2.0}

    def rhs_cfg(self, static_values, states):
        nx = int(static_values['nx'])
        dx = float(static_values['dx'])
        if float(static_values['zero_bc']) != 0.0 or float(static_values['pi_bc']) != 0.0:
            raise ValueError('only zero Dirichlet boundary values are implemented (pde_control.py:32-35)')
        return (dx, float(static_values['beta_t']), float(static_values['beta_u']), float(static_values['gamma']),
                float(round(0.3 * nx)), float(round(0.7 * nx)))


lotka_volterra = LotkaVolterra()
van_der_pol = VanDerPol()
lunar_ascent = LunarAscent()
affine = Affine()
quad_decay = QuadDecay()
general3 = General3()
diffusion_reaction = DiffusionReaction()


def _size(shape):
    return int(math.prod(shape)) if shape else 1
