"""The four approaches (``ozone/integrators/*.py``) on top of the device engine.

Each class mirrors its reference namesake: constructed by ``ODEProblem.solve`` with the same
arguments (``ozone/integrators/integrator.py:14-34``), ``integrate()`` runs the solve,
``finalize_outputs()`` collates ``IntegratorOutputs`` (``integrator.py:39-48``).
"""
from __future__ import annotations

import numpy as np
import torch

from . import engine
from .collections import IntegratorOutputs, Packing, States, DynamicParameters, _size
from .func_wrapper import FuncWrapper, ODEVars


def evaluate_device_rhs(rhs, ozone_vars: ODEVars, static_values: dict):
    """``f(ozone_vars)`` for a device RHS: evaluate at ``num_nodes`` nodes with oz_rhs_eval and fill
    d_states / profile_outputs / field_outputs.  The node axis is the batch axis of the kernel."""
    nn = ozone_vars.num_nodes
    sts, dps = rhs.names('states'), rhs.names('dynamic_parameters')
    ys = [ozone_vars.states[n].reshape(nn, -1).t() for n, _ in sts]          # KeyError if absent
    ps = [ozone_vars.dynamic_parameters[n].reshape(nn, -1).t() for n, _ in dps]
    y = torch.cat(ys, dim=0).contiguous()
    if not y.is_cuda:
        raise RuntimeError('device right-hand sides can only be evaluated on a CUDA device')
    p = torch.cat(ps, dim=0).contiguous() if ps else None
    th = None
    if rhs.static_parameters:
        from .utils import variablize
        th = torch.cat([variablize(static_values[n]).to(y).reshape(-1) for n, _ in rhs.names('static_parameters')])
    spec = engine.SolveSpec(rhs_name=rhs.kernel, method=_dummy_method(), N=nn, n=1, dtype=y.dtype,
                            th_shared=True, d_override=y.shape[0] if rhs.kind == 'grid' else 0,
                            rhs_cfg=rhs.rhs_cfg(static_values, None) if rhs.kind == 'grid' else ())
    dy, prof, fld = engine.rhs_eval(spec, y, p, th)
    off = 0
    for (n, _), blk in zip(sts, ys):
        k = blk.shape[0]
        ozone_vars.d_states[n] = dy[off:off + k].t().reshape(ozone_vars.states[n].shape)
        off += k
    for decl, arr, dst in ((rhs.names('profile_outputs'), prof, ozone_vars.profile_outputs),
                           (rhs.names('field_outputs'), fld, ozone_vars.field_outputs)):
        off = 0
        for n, shp in decl:
            k = _size(shp)
            dst[n] = arr[off:off + k].t().reshape((nn,) + (tuple(shp) if shp else (1,)))
            off += k


def default_newton_tolerance(dtype):
    """1e-7 in fp64: the value the reference hard-codes (time_marching.py:220).  The kernel stops on the 2-norm of the
    stacked stage residual Y - h A F - y, which in fp32 is quantised at ~1e-7 |y| per component, so 1e-7 could only be
    met by an exactly zero residual and every step would run into newton_max_iter; fp32 problems default to 1e-5."""
    return 1e-7 if dtype == torch.float64 else 1e-5


def _dummy_method():
    from .glms import get_integration_method
    return get_integration_method('ForwardEuler')


class IntegratorBase(object):
    def __init__(self, method, approach, ode_func: FuncWrapper, states: States, dynamic_parameters: DynamicParameters,
                 num_times: int, step_vector, options=None):
        self.method = method
        self.approach = approach
        self.ode_func = ode_func
        self.states = states
        self.dynamic_parameters = dynamic_parameters
        self.num_times = num_times
        self.num_steps = num_times - 1
        self.step_vector = step_vector
        self.options = dict(options or {})
        self.outputs = IntegratorOutputs()

    # ---- shared set-up -------------------------------------------------------------------
    def _prepare(self):
        f = self.ode_func
        if f.is_device:
            rhs = f.f
        else:
            rhs = self._trace_python_callback()                          # KeyError / ValueError / TypeError like the reference
        self.rhs = rhs
        static_values = rhs.bind_static(f.args, f.kwargs)               # TypeError like a wrong Python call
        self.packing = pk = Packing(rhs, self.states, self.dynamic_parameters, static_values, self.step_vector,
                                    self.num_times)
        if pk.device.type != 'cuda':
            raise RuntimeError('ozone_b200 solves on a CUDA device only (no CPU fallback); '
                               f'the states live on {pk.device}')
        self.y0 = pk.pack_states(self.states)
        self.h = pk.pack_h(self.step_vector)
        self.p = pk.pack_params(self.dynamic_parameters)
        self.th = pk.pack_static()
        opts = self.options
        traced = None if f.is_device else rhs
        self.spec = engine.SolveSpec(
            rhs_name='traced' if traced is not None else rhs.kernel, traced=traced, method=self.method, N=pk.N,
            n=self.num_steps, dtype=pk.dtype,
            h_shared=pk.h_shared, p_shared=pk.p_shared, th_shared=pk.th_shared,
            store_history=any(s.output for s in self.states.states.values()),
            newton_tol=opts.get('newton_tolerance', default_newton_tolerance(pk.dtype)),
            newton_max_iter=opts.get('newton_max_iter', 100),
            d_override=pk.d if pk.grid else 0, rhs_cfg=rhs.rhs_cfg(static_values, self.states),
            state_blocks=tuple(pk.state_offsets()))
        return rhs

    def _trace_python_callback(self):
        """Run the Python callback once on symbolic inputs (two nodes, so shapes are checked the way the
        reference checks them with num_nodes > 1) and turn it into a device functor (trace.py).  Malformed
        callbacks raise the reference's KeyError / ValueError / TypeError (ozone/func_wrapper.py:28-53)."""
        from .trace import TracedRHS, _norm_shape, MAX_STATE
        batched = bool(self.options.get('batched', False))
        f = self.ode_func
        st = self.states.states
        if not batched and len(st) == 1:
            # ONE state that is a long vector: a 1-D grid (pde_control.py:11-81), compiled into the fused stencil kernels
            (name, s), = st.items()
            shp = tuple(s.shape)
            if len(shp) == 1 and shp[0] > MAX_STATE:
                from .trace_grid import TracedGridRHS
                for pn, dp in self.dynamic_parameters.d_params.items():
                    if int(np.prod(tuple(dp.point_shape) or (1,))) != 1:
                        raise ValueError(f'dynamic parameter \'{pn}\' of a grid right-hand side must be one scalar per time point, '
                                         f'got point shape {tuple(dp.point_shape)}')
                return TracedGridRHS(f.f, f.args, f.kwargs, name, shp[0], list(self.dynamic_parameters.d_params))
        N = 1
        state_shapes, param_shapes = {}, {}
        for n, s in self.states.states.items():
            shp = tuple(s.shape)
            if batched:
                if len(shp) == 0:
                    raise ValueError(f'batched=True: state \'{n}\' needs a trailing trajectory axis')
                N, shp = max(N, shp[-1]), shp[:-1]
            state_shapes[n] = _norm_shape(shp)
        for n, d in self.dynamic_parameters.d_params.items():
            shp = tuple(d.point_shape)
            if batched and N > 1 and len(shp) >= 1 and shp[-1] == N:
                shp = shp[:-1]
            param_shapes[n] = _norm_shape(shp)
        f = self.ode_func
        return TracedRHS(f.f, f.args, f.kwargs, state_shapes, param_shapes, batch_size=N)

    def _collect(self, out):
        pk, rhs = self.packing, self.rhs
        nt = self.num_times
        res = self.outputs
        if out.get('hist') is not None:
            named = pk.unpack_states(out['hist'], self.states, (nt,))
            res.states = {n: v for n, v in named.items() if self.states.states[n].output}
        finals = pk.unpack_states(out['final'], self.states, ())
        res.final_states = {n: v for n, v in finals.items() if self.states.states[n].store_final}
        for n, s in self.states.states.items():
            s.solved_state = res.states.get(n)
            s.final_state = res.final_states.get(n)
        if out.get('prof') is not None:
            res.profile_outputs = pk.unpack_outputs(out['prof'], rhs.names('profile_outputs'), (nt,))
        if out.get('field') is not None:
            res.field_outputs = pk.unpack_outputs(out['field'], rhs.names('field_outputs'), ())
        res.status, res.iterations = out.get('status'), out.get('iters')

    def finalize_outputs(self) -> IntegratorOutputs:
        return self.outputs

    def integrate(self):
        raise NotImplementedError


class TimeMarching(IntegratorBase):
    """ozone/integrators/time_marching.py."""

    def integrate(self):
        self._prepare()
        self._collect(engine.time_march(self.spec, self.y0, self.h, self.p, self.th))


class TimeMarchingUniform(IntegratorBase):
    """ozone/integrators/time_marching_uniform.py: checkpointed reverse sweep."""

    def integrate(self):
        from dataclasses import replace
        num_checkpoints = self.approach.num_checkpoints
        if num_checkpoints is None:
            num_checkpoints = int(self.num_steps ** 0.5)                     # time_marching_uniform.py:19-21
        self.num_checkpoints = num_checkpoints
        if num_checkpoints <= 0 or not (self.num_steps % num_checkpoints == 0):
            raise ValueError(f'Number of timesteps ({self.num_steps}) must be divisible by number of checkpoints {num_checkpoints}')
        self.num_inner_steps = self.num_steps // num_checkpoints
        self._prepare()
        self.spec = replace(self.spec, num_inner=self.num_inner_steps)
        if self.options.get('verbose', True):                                # time_marching_uniform.py:34-40
            print('\nCheckpoint information:================')
            print('   # of points:       ', self.num_times)
            print('   # of steps:        ', self.num_steps)
            print('   # of checkpoints:  ', self.num_checkpoints)
            print('   # of inner iter:   ', self.num_inner_steps)
            print('   # of outer iter:   ', self.num_checkpoints)
            print('=======================================\n')
        self._collect(engine.time_march(self.spec, self.y0, self.h, self.p, self.th))


class PicardIteration(IntegratorBase):
    """ozone/integrators/picard_iteration.py (+ vector_based.py for the outputs)."""

    def integrate(self):
        self._prepare()
        tol = self.approach.tolerance
        if tol is None:
            tol = 1e-10 if self.spec.dtype == torch.float64 else 1e-4   # fp32 round-off floor of the residual 2-norm is ~1e-5
        guess = self.packing.pack_guess(self.states)
        if self.approach.relaxation == 'jacobi':
            from dataclasses import replace
            self.spec = replace(self.spec, state_blocks=(0, self.spec.d))
        self._collect(engine.picard(self.spec, self.y0, self.h, self.p, self.th, guess=guess, tol=tol,
                                    max_sweeps=self.approach.max_iter))


class Collocation(IntegratorBase):
    """ozone/integrators/collocation.py: the stage and state unknowns become design variables, the GLM
    equations equality constraints; nothing is solved here -- an optimiser drives them to zero."""

    def integrate(self):
        rhs = self._prepare()
        pk, n, s = self.packing, self.num_steps, self.method.num_stages
        guess = pk.pack_guess(self.states)
        if guess is None:                                                    # collections.py:129-133
            guess = self.y0.detach().unsqueeze(0).expand(n, pk.d, pk.N)
        given = self.options.get('collocation_design_values')
        if given is not None:            # an optimiser re-evaluating the model at its own iterate (optimize.py)
            stage_dv, state_dv = given
            if tuple(stage_dv.shape) != (n, s, pk.d, pk.N) or tuple(state_dv.shape) != (n, pk.d, pk.N):
                raise ValueError(f'collocation_design_values must have shapes {(n, s, pk.d, pk.N)} and {(n, pk.d, pk.N)}')
        else:
            stage_dv = guess.unsqueeze(1).expand(n, s, pk.d, pk.N).contiguous().clone().requires_grad_(True)
            state_dv = guess.contiguous().clone().requires_grad_(True)
        c1, c2 = engine.collocation(self.spec, self.y0, self.h, self.p, self.th, stage_dv, state_dv)
        hist = torch.cat([self.y0.unsqueeze(0), state_dv], dim=0)            # collections.py:228-232
        out = dict(hist=hist, final=hist[-1])
        if rhs.profile_outputs or rhs.field_outputs:                         # vector_based.py:33-49
            prof, fld = engine.outputs_on_history(self.spec, hist, self.p, self.th)
            out.update(prof=prof, field=fld)
        self.spec.store_history = True
        self._collect(dict(out, hist=hist))
        if not any(st.output for st in self.states.states.values()):
            self.outputs.states = {}
        # design variables / constraints, per state, with the reference's names and scalers
        self.design_variables, self.constraints, self.scalers = {}, {}, {}
        off = 0
        for name, ps, size, _ in pk.state_meta:
            st = self.states.states[name]
            g_state = guess[:, off:off + size]
            if st.scaler is None:                                            # collections.py:204-209
                m = float(g_state.mean())
                sc = 1.0 / float(g_state.abs().mean()) if abs(m) > 1e-7 else 1.0
                stage_sc = state_sc = sc
            else:
                stage_sc = state_sc = st.scaler
            self.design_variables[f'{name}_stage_dv'] = (stage_dv, slice(off, off + size))
            self.design_variables[f'{name}_state_dv'] = (state_dv, slice(off, off + size))
            self.constraints[f'{name}_stage_c'] = c1[:, :, off:off + size]
            self.constraints[f'{name}_state_c'] = c2[:, off:off + size]
            self.scalers[name] = dict(stage=stage_sc, state=state_sc)
            off += size
        self.stage_dv, self.state_dv, self.c1, self.c2 = stage_dv, state_dv, c1, c2
