"""Problem data containers (``ozone/collections.py``) and the packing between the user's named
tensors and the device layout (structure of arrays, trajectory index innermost).

Batch convention: a state whose functor point-shape is ``ps`` is given as a tensor of shape ``ps``
(one trajectory; a scalar is shape ``(1,)`` like a scalar csdl.Variable) or ``ps + (N,)`` (N
trajectories).  This is how the reference's test_3_large expresses 500 independent ODEs
(``test_3_large/utils.py:7-38``).  Dynamic parameters are ``(num_times,) + ps`` (shared by the
batch) or ``(num_times,) + ps + (N,)``; static parameters ``ps`` or ``ps + (N,)``.
"""
from __future__ import annotations

import math
from typing import Optional

import numpy as np
import torch


def _size(shape):
    return int(math.prod(shape)) if len(shape) else 1


class State(object):
    def __init__(self, name, initial_condition, output, store_final, initial_guess, scaler) -> None:
        self.name = name
        self.initial_condition = initial_condition
        self.output = output
        self.store_final = store_final
        self.initial_guess = initial_guess
        self.scaler = scaler
        self.shape = tuple(initial_condition.shape)
        self.size = initial_condition.numel()
        self.solved_shape = None
        self.solved_state = None
        self.final_state = None


class States(object):
    def __init__(self) -> None:
        self.states: dict[str, State] = {}
        self.full_size = 0

    def add(self, name, initial_condition, output, store_final, initial_guess, scaler):
        if name in self.states:
            raise ValueError(f'state with name \'{name}\' already exists')
        self.states[name] = State(name, initial_condition, output, store_final, initial_guess, scaler)

    def validate(self, num_times: int):
        """ozone/collections.py:84-95."""
        if len(self.states) == 0:
            raise ValueError('no states added to the ODE problem')
        self.full_size = 0
        for name, state in self.states.items():
            state.solved_shape = (num_times,) + state.shape
            self.full_size += state.size
            if state.initial_guess is not None:
                if tuple(state.initial_guess.shape) != (num_times - 1,) + state.shape:
                    raise ValueError(f'Expected shape {(num_times - 1,) + state.shape} for initial guess of state \'{name}\' but got {tuple(state.initial_guess.shape)}')


class DynamicParameter(object):
    def __init__(self, name, parameter) -> None:
        self.name = name
        self.parameter = parameter
        self.given_shape = tuple(parameter.shape)
        self.point_shape = tuple(parameter.shape[1:])


class DynamicParameters(object):
    def __init__(self) -> None:
        self.d_params: dict[str, DynamicParameter] = {}

    def add(self, name, parameter):
        if name in self.d_params:
            raise ValueError(f'dynamic parameter with name \'{name}\' already exists')
        if parameter.dim() == 1:                      # ozone/collections.py:247-248
            parameter = parameter.reshape((parameter.shape[0], 1))
        self.d_params[name] = DynamicParameter(name, parameter)

    def validate(self, num_times: int):
        for name, param in self.d_params.items():
            if param.parameter.shape[0] != num_times:
                raise ValueError(f'Expected first axis of dynamic parameter \'{name}\' to be size {num_times} but got {param.parameter.shape[0]}')


class IntegratorOutputs(object):
    """ozone/collections.py:335-355."""

    def __init__(self) -> None:
        self.states = {}
        self.final_states = {}
        self.profile_outputs = {}
        self.field_outputs = {}
        # extras of the device engine (not in the reference): per-trajectory solver status
        self.status = None            # int32 (N,): bit 1 not converged, 2 singular, 4 non-finite; None if all clear is unknown
        self.iterations = None        # int32 (N,): Newton iterations (implicit time marching) / Picard sweeps

    def __str__(self):
        groups = [('States:', self.states), ('Final States:', self.final_states),
                  ('Profile Outputs:', self.profile_outputs), ('Field Outputs:', self.field_outputs)]
        string = '\nIntegrator outputs:\n'
        for title, vars_ in groups:
            string += f'  {title}\n'
            for key, value in vars_.items():
                string += f'    {key}: {tuple(value.shape)}\n'
        return string


# ------------------------------------------------------------------------------------------------
class Packing:
    """Maps named user tensors <-> packed device arrays for one (problem, device RHS) pair."""

    def __init__(self, rhs, states: States, d_params: DynamicParameters, static_values: dict, step_vector,
                 num_times: int):
        self.rhs = rhs
        self.num_times = num_times
        decl_states = rhs.names('states')
        decl_dp = rhs.names('dynamic_parameters')
        # --- name matching: what indexing ozone_vars.states[name] / setting d_states[name] would do
        for n, _ in decl_states:
            if n not in states.states:
                raise KeyError(f'the ODE function reads state \'{n}\' but it was not added to the problem. Given states: {list(states.states)}')
        for n in states.states:
            if n not in dict(decl_states):
                raise KeyError(f'No state derivative specified for state \'{n}\'.')
        for n, _ in decl_dp:
            if n not in d_params.d_params:
                raise KeyError(f'the ODE function reads dynamic parameter \'{n}\' but it was not added. Given: {list(d_params.d_params)}')

        ref = next(iter(states.states.values())).initial_condition
        self.device, self.dtype = ref.device, ref.dtype
        if self.dtype not in (torch.float64, torch.float32):
            raise TypeError(f'states must be float64 or float32 tensors, got {self.dtype}')

        # --- batch size
        self.grid = rhs.kind == 'grid'
        self.state_meta = []          # (name, point_shape, size, batched)
        N = 1
        for n, ps in decl_states:
            ic = states.states[n].initial_condition
            shp = tuple(ic.shape)
            if self.grid:
                if len(shp) != 1:
                    raise ValueError(f'grid state \'{n}\' must be a 1-D tensor of cells, got shape {shp}')
                self.state_meta.append((n, shp, shp[0], False))
                continue
            nb = self._batch_of(shp, ps, f'state \'{n}\'')
            self.state_meta.append((n, ps, _size(ps), nb is not None))
            if nb is not None:
                N = self._merge(N, nb, n)
        self.dp_meta = []
        for n, ps in decl_dp:
            par = d_params.d_params[n].parameter
            shp = tuple(par.shape[1:])
            nb = None if self.grid else self._batch_of(shp, ps, f'dynamic parameter \'{n}\'')
            if self.grid and _size(shp) != 1:
                raise ValueError(f'dynamic parameter \'{n}\' must be scalar per time point, got {shp}')
            self.dp_meta.append((n, ps, _size(ps), nb is not None))
            if nb is not None:
                N = self._merge(N, nb, n)
        self.st_meta = []
        self.static = {}
        for n, ps in rhs.names('static_parameters'):
            from .utils import variablize
            v = variablize(static_values[n], n).to(device=self.device, dtype=self.dtype)
            nb = self._batch_of(tuple(v.shape), ps, f'static parameter \'{n}\'')
            self.st_meta.append((n, ps, _size(ps), nb is not None))
            self.static[n] = v
            if nb is not None:
                N = self._merge(N, nb, n)
        self.N = N
        self.d = sum(m[2] for m in self.state_meta)
        self.P = sum(m[2] for m in self.dp_meta)
        self.S = sum(m[2] for m in self.st_meta)
        self.p_shared = not any(m[3] for m in self.dp_meta)
        self.th_shared = not any(m[3] for m in self.st_meta)
        h = step_vector
        self.h_shared = h.dim() == 1
        if not self.h_shared and h.shape[1] != N:
            raise ValueError(f'per-trajectory step_vector must have shape ({num_times - 1}, {N}), got {tuple(h.shape)}')

    @staticmethod
    def _batch_of(shape, ps, what) -> Optional[int]:
        """None if `shape` is one point, else the batch size on the trailing axis."""
        if shape == tuple(ps) or (len(ps) == 0 and shape in ((), (1,))):
            return None
        if len(shape) == len(ps) + 1 and tuple(shape[:-1]) == tuple(ps):
            return int(shape[-1])
        raise ValueError(f'Expected shape {tuple(ps)} or {tuple(ps) + ("N",)} for {what} but got {shape}')

    @staticmethod
    def _merge(N, nb, name):
        if N != 1 and nb != 1 and nb != N:
            raise ValueError(f'inconsistent batch sizes: \'{name}\' has {nb} trajectories, others have {N}')
        return max(N, nb)

    def state_offsets(self):
        """Packed-component offsets [0, size_0, size_0 + size_1, ..., d] of the states in insertion order."""
        offs = [0]
        for m in self.state_meta:
            offs.append(offs[-1] + m[2])
        return offs if (not self.grid and len(offs) <= 9) else []

    # --- inputs -> packed (differentiable torch ops: cat / reshape / expand)
    def pack_states(self, states: States):
        N = self.N
        rows = []
        for n, ps, size, batched in self.state_meta:
            ic = states.states[n].initial_condition.to(self.dtype)
            rows.append(ic.reshape(size, -1).expand(size, N))
        return torch.cat(rows, dim=0).contiguous()

    def pack_guess(self, states: States):
        """collections.py:125-133: user guess (n, *shape) per state, else None (= broadcast IC)."""
        if all(s.initial_guess is None for s in states.states.values()):
            return None
        n = self.num_times - 1
        rows = []
        for name, ps, size, batched in self.state_meta:
            st = states.states[name]
            if st.initial_guess is None:
                g = st.initial_condition.to(self.dtype).reshape(1, size, -1).expand(n, size, self.N)
            else:
                g = torch.as_tensor(st.initial_guess, dtype=self.dtype, device=self.device).reshape(n, size, -1).expand(n, size, self.N)
            rows.append(g)
        return torch.cat(rows, dim=1).contiguous()

    def pack_params(self, d_params: DynamicParameters):
        if self.P == 0:
            return None
        nt, N = self.num_times, self.N
        rows = []
        for n, ps, size, batched in self.dp_meta:
            par = d_params.d_params[n].parameter.to(self.dtype)
            if self.p_shared:
                rows.append(par.reshape(nt, size))
            else:
                rows.append(par.reshape(nt, size, -1).expand(nt, size, N))
        return torch.cat(rows, dim=1).contiguous()

    def pack_static(self):
        if self.S == 0:
            return None
        rows = []
        for n, ps, size, batched in self.st_meta:
            v = self.static[n]
            rows.append(v.reshape(size) if self.th_shared else v.reshape(size, -1).expand(size, self.N))
        return torch.cat(rows, dim=0).contiguous()

    def pack_h(self, step_vector):
        return step_vector.to(self.dtype).contiguous()

    # --- packed -> named outputs (views, no copies)
    def _point(self, ps, batched_out):
        ps = tuple(ps)
        if batched_out:
            return ps + (self.N,)
        return ps if len(ps) else (1,)

    def unpack_states(self, hist, states: States, lead):
        """hist (nt, d, N) or (d, N) -> {name: (lead..., *ic.shape or *ps, N)}"""
        out, off = {}, 0
        for n, ps, size, batched in self.state_meta:
            blk = hist[..., off:off + size, :]
            off += size
            if self.grid:
                out[n] = blk.reshape(lead + (size,))
            else:
                out[n] = blk.reshape(lead + self._point(ps, self.N > 1))
        return out

    def unpack_outputs(self, arr, decl, lead):
        out, off = {}, 0
        for n, ps in decl:
            size = _size(ps)
            blk = arr[..., off:off + size, :]
            off += size
            out[n] = blk.reshape(lead + self._point(ps, self.N > 1))
        return out
