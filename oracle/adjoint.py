"""Oracle derivatives (TEST INFRASTRUCTURE, see oracle/__init__.py).

The reference has no adjoint source: derivatives are requested from csdl_alpha
(``examples/simple_examples/lotka_volterra.py:76-79``, ``ozone/tests/run_ode.py:66-69``), which
reverse-differentiates the recorded time loop (``time_marching.py:51`` stack_all=True) and the
Newton implicit operation.  What is pinned are the resulting numbers (Jacobian norms in the
reference tests).  Two independent restatements are kept here and checked against each other
and against those goldens:
  * ``tm_adjoint``      hand-written discrete adjoint of the GLM step (what the CUDA kernels do);
  * ``complex_step``    forward complex-step differentiation of the whole solve (explicit methods).
"""
import numpy as np

from .integrators import (_as_method, _norm_inputs, _hk, interpolate_params, explicit_stages,
                          implicit_stages)


def _sum_to(x, shape):
    """Reduce a broadcast gradient back to the shape of a shared input."""
    x = np.asarray(x)
    while x.ndim > len(shape):
        x = x.sum(axis=-1)
    for ax, (a, b) in enumerate(zip(x.shape, shape)):
        if b == 1 and a != 1:
            x = x.sum(axis=ax, keepdims=True)
    return x.reshape(shape)


def tm_adjoint(rhs, method, y0, h, p=None, th=None, history=None, hist_bar=None, final_bar=None,
               prof_bar=None, field_bar=None, tol=1e-7, max_iter=100):
    """Discrete adjoint of time_marching for L = <hist_bar, history> + <final_bar, y_n>
    + <prof_bar, profile> + <field_bar, field>.  Returns gradients shaped like the inputs
    (shared inputs receive the sum over trajectories).

    Per step (lam = dL/dy_{k+1}):
        G_i  = h b_i lam + h sum_j a_ji Ybar_j ,  Ybar_j = J_y(Y_j)^T G_j
               (explicit: back-substitution i = s-1..0; implicit: the (s d) linear system
                M^T Ybar = blk(J^T) (h b (x) lam) with M = I - h (A (x) I) blk(J), the Newton matrix)
        pbar_i = J_p(Y_i)^T G_i -> p[k] (1-c_i), p[k+1] c_i   (transpose of collections.py:277-289)
        hbar_k = sum_i b_i F_i.lam + sum_i Ybar_i . sum_j a_ij F_j
        lam    <- lam + sum_i Ybar_i  (+ direct cotangents of y_k)
    """
    A, b, c, explicit = _as_method(method)
    y0, h, p, th, n = _norm_inputs(rhs, y0, h, p, th)
    d, N = y0.shape
    s = A.shape[0]
    if history is None:
        from .integrators import time_marching
        history = time_marching(rhs, (A, b, c, explicit), y0, h, p, th, tol, max_iter)['history']
    ps = interpolate_params(p, c)
    zero = lambda shp: np.zeros(shp)
    hist_bar = zero(history.shape) if hist_bar is None else np.broadcast_to(hist_bar, history.shape)
    final_bar = zero((d, N)) if final_bar is None else np.broadcast_to(final_bar, (d, N))
    prof_bar = zero((n + 1, rhs.NPROF, N)) if prof_bar is None else np.broadcast_to(prof_bar, (n + 1, rhs.NPROF, N))
    field_bar = zero((rhs.NFIELD, N)) if field_bar is None else np.broadcast_to(field_bar, (rhs.NFIELD, N))

    pb = np.zeros((n + 1, rhs.P, N))
    tb = np.zeros((rhs.S, N))
    hb = np.zeros((n, N))
    has_out = rhs.NPROF + rhs.NFIELD > 0

    def direct(k):
        g = hist_bar[k].copy()
        if has_out:
            yb, pbk, tbk = rhs.out_vjp(history[k], p[k], th, prof_bar[k], field_bar)
            g = g + yb
            pb[k] += np.broadcast_to(pbk, (rhs.P, N)) if rhs.P else 0
            if rhs.S:
                tb[:] += np.broadcast_to(tbk, (rhs.S, N))
        return g

    lam = final_bar + direct(n)
    for k in range(n - 1, -1, -1):
        y = history[k]
        hk = _hk(h, k)
        if explicit:
            Y, F = explicit_stages(rhs, A, y, hk, ps[k], th)
        else:
            Y, F, _, _ = implicit_stages(rhs, A, y, hk, ps[k], th, tol, max_iter)
        Ybar = np.zeros_like(Y)
        G = np.zeros_like(Y)
        if explicit:
            for i in range(s - 1, -1, -1):
                G[i] = hk * b[i] * lam + hk * np.einsum('j,jdn->dn', A[:, i], Ybar)
                Ybar[i] = rhs.vjp(Y[i], ps[k][i], th, G[i])[0]
        else:
            m = s * d
            hkN = np.broadcast_to(np.asarray(hk, dtype=float), (N,))
            J = np.stack([np.broadcast_to(rhs.jac(Y[j], ps[k][j], th), (d, d, N)) for j in range(s)])   # (s, d, d, N)
            M = np.zeros((N, s, d, s, d))
            for i in range(s):
                for j in range(s):
                    M[:, i, :, j, :] -= (hkN * A[i, j])[:, None, None] * np.moveaxis(J[j], -1, 0)
            M = M.reshape(N, m, m) + np.eye(m)
            rhsv = np.stack([np.einsum('ijn,in->jn', J[i], hkN * b[i] * lam) for i in range(s)])        # (s, d, N)
            sol = np.linalg.solve(np.swapaxes(M, 1, 2), np.moveaxis(rhsv.reshape(m, N), -1, 0)[..., None])[..., 0]
            Ybar = np.moveaxis(sol, 0, -1).reshape(s, d, N)
            for i in range(s):
                G[i] = hk * b[i] * lam + hk * np.einsum('j,jdn->dn', A[:, i], Ybar)
        for i in range(s):
            _, pbi, tbi = rhs.vjp(Y[i], ps[k][i], th, G[i])
            if rhs.P:
                pbi = np.broadcast_to(pbi, (rhs.P, N))
                pb[k] += (1.0 - c[i]) * pbi
                pb[k + 1] += c[i] * pbi
            if rhs.S:
                tb += np.broadcast_to(tbi, (rhs.S, N))
        AF = np.einsum('ij,jdn->idn', A, F)
        hb[k] = np.einsum('i,idn,dn->n', b, F, lam) + np.einsum('idn,idn->n', Ybar, AF)
        lam = lam + Ybar.sum(axis=0) + direct(k)
    return dict(y0=lam, h=_sum_to(hb, h.shape), p=_sum_to(pb, p.shape), th=_sum_to(tb, th.shape))


def complex_step(solve, inputs, wrt, eps=1e-30):
    """d(scalar solve(**inputs))/d(inputs[wrt]) by complex step, entry by entry (small cases)."""
    base = np.asarray(inputs[wrt], dtype=float)
    g = np.zeros(base.shape)
    for idx in np.ndindex(*base.shape):
        xp = base.astype(complex)
        xp[idx] += 1j * eps
        kw = dict(inputs)
        kw[wrt] = xp
        g[idx] = np.imag(solve(**kw)) / eps
    return g


def colloc_vjp(rhs, method, y0, h, Ydv, ydv, p=None, th=None, mu1=None, mu2=None):
    """Gradient of <mu1, c1> + <mu2, c2> (ozone/integrators/collocation.py:48-50) with respect to
    Ydv, ydv, y0, h, p, th -- what an optimiser consumes as (dc)^T mu."""
    A, b, c, _ = _as_method(method)
    y0, h, p, th, n = _norm_inputs(rhs, y0, h, p, th)
    d, N = y0.shape
    s = A.shape[0]
    ps = interpolate_params(p, c)
    hN = np.broadcast_to(h if h.ndim == 2 else h[:, None], (n, N))
    # cotangent on hF_kj : sum_i a_ij mu1_ki + b_j mu2_k
    W = np.einsum('ij,kidn->kjdn', A, mu1) + b.reshape(1, s, 1, 1) * mu2[:, None]
    Yb = -mu1.copy()
    pb = np.zeros((n + 1, rhs.P, N))
    tb = np.zeros((rhs.S, N))
    hb = np.zeros((n, N))
    for k in range(n):
        for j in range(s):
            F = rhs.f(Ydv[k, j], ps[k][j], th)
            G = hN[k] * W[k, j]
            yb, pbj, tbj = rhs.vjp(Ydv[k, j], ps[k][j], th, G)
            Yb[k, j] += yb
            if rhs.P:
                pbj = np.broadcast_to(pbj, (rhs.P, N))
                pb[k] += (1 - c[j]) * pbj
                pb[k + 1] += c[j] * pbj
            if rhs.S:
                tb += np.broadcast_to(tbj, (rhs.S, N))
            hb[k] += np.einsum('dn,dn->n', np.broadcast_to(F, (d, N)), W[k, j])
    # yfull[k] enters c1_k (+), c2_k (+), c2_{k-1} (-)
    yfb = np.zeros((n + 1, d, N))
    yfb[:-1] += mu1.sum(axis=1) + mu2
    yfb[1:] -= mu2
    return dict(Y=Yb, y=yfb[1:], y0=yfb[0], h=_sum_to(hb, h.shape), p=_sum_to(pb, p.shape),
                th=_sum_to(tb, th.shape))
