"""Optimiser adapter for the Collocation approach.

In the reference, ``Collocation.solve`` only DECLARES design variables and equality constraints
(``ozone/integrators/collocation.py:8-57``); an external optimiser then asks the simulator for
``f, c, df, dc`` (``ozone/tests/run_ode.py:49-58``: modopt's SLSQP, ftol 1e-9, on a dummy objective).
modopt is not available here, scipy is, and scipy's SLSQP is the same algorithm (Kraft's code), so this
module exposes the problem in the ``scipy.optimize.minimize`` vocabulary:

    x    scaled design variables  [stage_dv * stage_scaler, state_dv * state_scaler]  (collections.py:204-225)
    c    scaled constraints       [c1 * stage_scaler, c2 * stage_scaler]               (collocation.py:56-57)
    dc   dense Jacobian, assembled by ONE batched launch of the collocation VJP kernel: the unit
         multipliers e_1..e_m are laid out as a batch of m trajectories (the kernel's natural axis)

The optimiser itself (dense SQP linear algebra on the host) is outside the hot path, as in the reference.
"""
from __future__ import annotations

import numpy as np
import torch

from . import engine
from .integrators import Collocation


class CollocationNLP:
    """Values and derivatives of the collocation equality constraints of a solved ``ODEProblem`` (N = 1)."""

    def __init__(self, ode_problem):
        integ = getattr(ode_problem, 'integrator', None)
        if not isinstance(integ, Collocation):
            raise TypeError('expected an ODEProblem solved with approaches.Collocation()')
        if integ.packing.N != 1:
            raise ValueError('the optimiser adapter handles one trajectory; batches are independent problems')
        self.integ = integ
        spec = integ.spec
        self.n, self.s, self.d = spec.n, spec.method.num_stages, spec.d
        self.nY, self.ny = self.n * self.s * self.d, self.n * self.d
        self.m = self.nY + self.ny
        dev, dt = integ.y0.device, spec.dtype
        # per-component scalers in packed order (every state has its own, collections.py:204-209)
        sc_stage = torch.ones(self.d, dtype=dt, device=dev)
        sc_state = torch.ones(self.d, dtype=dt, device=dev)
        off = 0
        for name, ps, size, _ in integ.packing.state_meta:
            sc_stage[off:off + size] = float(np.mean(integ.scalers[name]['stage']))
            sc_state[off:off + size] = float(np.mean(integ.scalers[name]['state']))
            off += size
        self.scale_x = torch.cat([sc_stage.repeat(self.n * self.s), sc_state.repeat(self.n)])
        self.scale_c = torch.cat([sc_stage.repeat(self.n * self.s), sc_stage.repeat(self.n)])   # both constraints: stage_scaler
        self.inputs = tuple(t.detach() if t is not None else None for t in (integ.y0, integ.h, integ.p, integ.th))
        # the Jacobian launch: m unit multipliers as a batch of m trajectories
        from dataclasses import replace
        self.jspec = replace(spec, N=self.m, h_shared=True, p_shared=spec.p_shared, th_shared=spec.th_shared)
        eye = torch.eye(self.m, dtype=dt, device=dev)
        self.mu1 = eye[:self.nY].reshape(self.n, self.s, self.d, self.m).contiguous()
        self.mu2 = eye[self.nY:].reshape(self.n, self.d, self.m).contiguous()

    # ---- packing of the unknowns
    def x0(self):
        z = torch.cat([self.integ.stage_dv.detach().reshape(-1), self.integ.state_dv.detach().reshape(-1)])
        return (z * self.scale_x).cpu().numpy()

    def _split(self, x):
        z = torch.as_tensor(x, dtype=self.scale_x.dtype, device=self.scale_x.device) / self.scale_x
        Y = z[:self.nY].reshape(self.n, self.s, self.d, 1).contiguous()
        y = z[self.nY:].reshape(self.n, self.d, 1).contiguous()
        return Y, y

    def constraints(self, x):
        Y, y = self._split(x)
        c1, c2 = engine.colloc_residual(self.integ.spec, *self.inputs, Y, y)
        return (torch.cat([c1.reshape(-1), c2.reshape(-1)]) * self.scale_c).cpu().numpy()

    def jacobian(self, x):
        """dense d(scaled c)/d(scaled x), shape (m, m)."""
        Y, y = self._split(x)
        y0, h, p, th = self.inputs
        m = self.m
        ex = lambda t, shared: t if (t is None or shared) else t.expand(*t.shape[:-1], m).contiguous()
        g = engine.colloc_vjp(self.jspec, ex(y0, False), h, ex(p, self.jspec.p_shared), ex(th, self.jspec.th_shared),
                              Y.expand(self.n, self.s, self.d, m).contiguous(), y.expand(self.n, self.d, m).contiguous(),
                              self.mu1, self.mu2, need=())
        J = torch.cat([g['Y'].reshape(self.nY, m), g['y'].reshape(self.ny, m)], dim=0).t()     # rows = constraints
        return (J * self.scale_c[:, None] / self.scale_x[None, :]).cpu().numpy()

    def commit(self, x):
        """Write a solution back into the integrator's design variables and rebuild its outputs."""
        Y, y = self._split(x)
        integ = self.integ
        with torch.no_grad():
            integ.stage_dv.copy_(Y)
            integ.state_dv.copy_(y)
        hist = torch.cat([integ.y0.detach().unsqueeze(0), integ.state_dv.detach()], dim=0)
        out = dict(hist=hist, final=hist[-1])
        if integ.rhs.profile_outputs or integ.rhs.field_outputs:
            prof, fld = engine.outputs_on_history(integ.spec, hist, self.inputs[2], self.inputs[3])
            out.update(prof=prof, field=fld)
        keep = {n for n, s in integ.states.states.items() if s.output}
        integ._collect(out)
        integ.outputs.states = {n: v for n, v in integ.outputs.states.items() if n in keep}
        return integ.outputs


def solve_collocation(ode_problem, ftol=1e-9, maxiter=200, verbose=False):
    """Drive the collocation constraints to zero with SLSQP on a dummy objective, exactly what the reference's
    test harness does (``ozone/tests/run_ode.py:24-33, 49-58``).  Returns the updated ``IntegratorOutputs``."""
    from scipy.optimize import minimize
    nlp = CollocationNLP(ode_problem)
    m = nlp.m
    # dummy design variable x_dummy = 2 with objective (x_dummy / 2)^2   (run_ode.py:29-32)
    x0 = np.concatenate([nlp.x0(), [2.0]])
    res = minimize(lambda z: (z[-1] / 2.0) ** 2, x0, jac=lambda z: np.concatenate([np.zeros(m), [z[-1] / 2.0]]),
                   method='SLSQP',
                   constraints=[{'type': 'eq', 'fun': lambda z: nlp.constraints(z[:-1]),
                                 'jac': lambda z: np.hstack([nlp.jacobian(z[:-1]), np.zeros((m, 1))])}],
                   options={'ftol': ftol, 'maxiter': maxiter, 'disp': verbose})
    outputs = nlp.commit(res.x[:-1])
    outputs.optimizer_result = res
    return outputs
