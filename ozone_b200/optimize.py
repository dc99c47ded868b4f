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


# ================================================================================================
# Batched {f, c, df, dc} for an optimiser (what `compute_optimization_derivatives()` hands modopt /
# SNOPT in the reference, ozone/paper_examples/utils/utils.py:237, 273-286).
class CollocationBlocks:
    """The Jacobian of the collocation constraints of N independent designs in factored, per-step form.

    For step k, stage i:  c1[k,i] = h_k sum_j a_ij f(Y[k,j], p_kj) + yfull[k] - Y[k,i]
                          c2[k]   = h_k sum_j b_j  f(Y[k,j], p_kj) + yfull[k] - yfull[k+1]          (collocation.py:48-50)
    with yfull = [y0; y] and p_kj = p[k] (1 - c_j) + p[k+1] c_j (collections.py:277-289).  Every block of dc is a product
    of a tableau entry, a step size and a node Jacobian, so only those are stored (one kernel launch, HBM-lean):
        F     [n, s, d, N]        f at the stage nodes                               (d c / d h_k)
        Jy_nz [n, s, nnz_y, N]    possibly-nonzero entries of df/dy (pattern_y, row-major)  (d c / d Y)
        Jp_nz [n, s, nnz_p, N]    possibly-nonzero entries of df/dp (pattern_p)             (d c / d p[k], d p[k+1])
    (`Jy`, `Jp` expand them to dense [n, s, d, d | P, N] on demand.)
    ``dense(i)`` expands one design's Jacobian; ``rmatvec`` is the transposed product through the VJP kernel."""

    def __init__(self, integ, Y, y):
        spec = integ.spec
        self.integ, self.spec = integ, spec
        n, s, d, N = spec.n, spec.method.num_stages, spec.d, spec.N
        P = spec.info['P']
        self.n, self.s, self.d, self.N, self.P = n, s, d, N, P
        y0, h, p, th = (None if t is None else t.detach() for t in (integ.y0, integ.h, integ.p, integ.th))
        A = torch.as_tensor(spec.method.A, dtype=spec.dtype, device=Y.device)
        B = torch.as_tensor(spec.method.B, dtype=spec.dtype, device=Y.device).reshape(-1)
        Cc = torch.as_tensor(spec.method.C, dtype=spec.dtype, device=Y.device).reshape(-1)
        self.A, self.B, self.C = A, B, Cc
        self.h = h if not spec.h_shared else h.unsqueeze(-1).expand(n, N)
        # ONE launch over the stage variables as they lie in memory; only the entries the right-hand side's sparsity
        # patterns allow leave the chip (lunar ascent: 9 of 49 and 3 of 21)
        self.pattern_y, self.pattern_p = engine.rhs_jac_pattern(spec, engine.KIND_COLLOC)
        F, Jy, Jp = engine.rhs_jac_eval(spec, Y.detach().reshape(n * s, d, N), p, th, compact=True, kind=engine.KIND_COLLOC,
                                        stage_params=True)       # the kernel interpolates p to the stage abscissae
        self.F = F.reshape(n, s, d, N)
        self.Jy_nz = Jy.reshape(n, s, -1, N)                     # nonzeros of df/dy in row-major order
        self.Jp_nz = Jp.reshape(n, s, -1, N) if Jp is not None else None

    def _expand(self, nz, pattern):
        """(n, s, nnz, N) -> (n, s, rows, cols, N) with the structural zeros filled in"""
        n, s, _, N = nz.shape
        full = torch.zeros((n, s) + pattern.shape + (N,), dtype=nz.dtype, device=nz.device)
        idx = torch.as_tensor(np.argwhere(pattern), device=nz.device)
        full[:, :, idx[:, 0], idx[:, 1], :] = nz
        return full

    @property
    def Jy(self):
        return self._expand(self.Jy_nz, self.pattern_y)

    @property
    def Jp(self):
        return self._expand(self.Jp_nz, self.pattern_p) if self.Jp_nz is not None else None

    # ---- one design, dense: rows [c1 (n s d), c2 (n d)], columns [Y (n s d), y (n d), y0 (d), h (n), p ((n+1) P)]
    def dense(self, i):
        n, s, d, P = self.n, self.s, self.d, self.P
        A, B, Cc = (t.cpu().numpy() for t in (self.A, self.B, self.C))
        F = self.F[..., i].cpu().numpy()
        Jy = self._expand(self.Jy_nz[..., i:i + 1], self.pattern_y)[..., 0].cpu().numpy()
        Jp = self._expand(self.Jp_nz[..., i:i + 1], self.pattern_p)[..., 0].cpu().numpy() if self.Jp_nz is not None else None
        h = self.h[:, i].cpu().numpy()
        nY, ny = n * s * d, n * d
        cols = dict(Y=(0, nY), y=(nY, ny), y0=(nY + ny, d), h=(nY + ny + d, n), p=(nY + ny + d + n, (n + 1) * P))
        ncol = nY + ny + d + n + (n + 1) * P
        J = np.zeros((nY + ny, ncol))
        yf_col = lambda k: (cols['y0'][0] if k == 0 else cols['y'][0] + (k - 1) * d)
        for k in range(n):
            for i_ in range(s):
                r0 = (k * s + i_) * d
                for j in range(s):
                    c0 = (k * s + j) * d
                    J[r0:r0 + d, c0:c0 + d] += h[k] * A[i_, j] * Jy[k, j]
                    if Jp is not None:
                        for q, wgt in ((k, 1.0 - Cc[j]), (k + 1, Cc[j])):
                            pc = cols['p'][0] + q * P
                            J[r0:r0 + d, pc:pc + P] += h[k] * A[i_, j] * wgt * Jp[k, j]
                    J[r0:r0 + d, cols['h'][0] + k] += A[i_, j] * F[k, j]
                J[r0:r0 + d, r0:r0 + d] -= np.eye(d)
                J[r0:r0 + d, yf_col(k):yf_col(k) + d] += np.eye(d)
            r0 = nY + k * d
            for j in range(s):
                c0 = (k * s + j) * d
                J[r0:r0 + d, c0:c0 + d] += h[k] * B[j] * Jy[k, j]
                if Jp is not None:
                    for q, wgt in ((k, 1.0 - Cc[j]), (k + 1, Cc[j])):
                        pc = cols['p'][0] + q * P
                        J[r0:r0 + d, pc:pc + P] += h[k] * B[j] * wgt * Jp[k, j]
                J[r0:r0 + d, cols['h'][0] + k] += B[j] * F[k, j]
            J[r0:r0 + d, yf_col(k):yf_col(k) + d] += np.eye(d)
            J[r0:r0 + d, yf_col(k + 1):yf_col(k + 1) + d] -= np.eye(d)
        return J, cols


class CollocationProblem:
    """``{f, c, df, dc}`` of a model built with ``approaches.Collocation()`` for a batch of N independent designs.

    ``model(dv)`` receives a dict of leaf tensors (the user's design variables, trajectory axis last) and the
    keyword ``collocation_design_values`` to pass to ``ODEProblem(...)``; it returns ``(ode_problem, objective,
    constraints)`` with ``objective`` of shape (N,) and ``constraints`` a dict name -> tensor whose LAST axis is the
    design axis.  This is the reference's ``build_recorder`` pattern (ascent_system/main_model.py:51-125) with
    ``set_as_design_variable / set_as_objective / set_as_constraint`` replaced by the return value.

    Scalers follow the reference: stage / state design variables and BOTH collocation constraints use
    ``1 / mean|guess|`` (collections.py:204-209, collocation.py:56-57); user variables use the scaler given here."""

    def __init__(self, model, design_variables: dict):
        self.model = model
        self.dv0 = {k: (v if isinstance(v, dict) else dict(value=v)) for k, v in design_variables.items()}
        self.Y = self.y = None

    def _leaves(self, values=None):
        out = {}
        for k, meta in self.dv0.items():
            v = meta['value'] if values is None or k not in values else values[k]
            out[k] = v.detach().clone().requires_grad_(True)
        return out

    def compute_optimization_derivatives(self, values=None, stage_dv=None, state_dv=None, dense_designs=()):
        """-> dict(f (N,), c {name: tensor}, df {dv: tensor}, dc {'user': {cname: {dv: tensor}}, 'collocation':
        CollocationBlocks}, scalers, x names).  User-constraint rows are differentiated one per-design row at a time
        (designs are independent, so one backward pass yields that row for all N designs)."""
        dv = self._leaves(values)
        given = None
        if stage_dv is not None:
            given = (stage_dv.detach().clone().requires_grad_(True), state_dv.detach().clone().requires_grad_(True))
        prob, f, cons = self.model(dv, collocation_design_values=given)
        integ = prob.integrator
        Y, y = integ.stage_dv, integ.state_dv
        self.Y, self.y, self.integ = Y, y, integ
        leaves = dict(dv, stage_dv=Y, state_dv=y)
        names = list(leaves)
        grads = torch.autograd.grad(f.sum(), [leaves[k] for k in names], retain_graph=True, allow_unused=True)
        df = {k: (g if g is not None else torch.zeros_like(leaves[k])) for k, g in zip(names, grads)}
        dc_user = {}
        for cname, c in cons.items():
            rows = c.reshape(-1, c.shape[-1])
            per = {k: [] for k in names}
            for r in range(rows.shape[0]):
                g = torch.autograd.grad(rows[r].sum(), [leaves[k] for k in names], retain_graph=True, allow_unused=True)
                for k, gi in zip(names, g):
                    per[k].append(gi)
            dc_user[cname] = {k: (torch.stack([gi if gi is not None else torch.zeros_like(leaves[k]) for gi in v])
                                  if any(gi is not None for gi in v) else None) for k, v in per.items()}
        blocks = CollocationBlocks(integ, Y, y)
        sc = integ.scalers
        cvals = dict(cons)
        cvals['stage_c'], cvals['state_c'] = integ.c1, integ.c2
        return dict(f=f.detach(), c={k: v.detach() for k, v in cvals.items()}, df=df,
                    dc=dict(user=dc_user, collocation=blocks), scalers=sc, design_variables=names,
                    problem=prob)
