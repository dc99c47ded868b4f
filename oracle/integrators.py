"""Oracle integrators (TEST INFRASTRUCTURE, see oracle/__init__.py): numpy fp64 restatement of
``ozone/integrators/*.py``.  Shapes: y (d, N); h (n,) or (n, N); p (nt, P, N|1); th (S, N|1).
"""
import numpy as np

from .tableaux import tableau


# --------------------------------------------------------------------------------------------
def interpolate_params(p, c):
    """ozone/collections.py:256-293: stage value p[k]*(1-c_i) + p[k+1]*c_i -> (n, s, P, N|1)."""
    p0, p1 = p[:-1, None], p[1:, None]
    cc = np.asarray(c).reshape(1, -1, 1, 1)
    return p0 * (1.0 - cc) + p1 * cc


def _hk(h, k):
    return h[k]          # scalar (shared) or (N,) (per trajectory); both broadcast against (d, N)


def explicit_stages(rhs, A, y, hk, pk, th):
    """ozone/integrators/time_marching.py:154-188.  Y_i = U_i*y + sum_j (hA)_ij F_j with the
    not-yet-computed F_j equal to zero (:156, :171-172); F_i = f(Y_i, p_i).  Returns Y, F of
    shape (s, d, N)."""
    s = A.shape[0]
    dt = np.result_type(y, np.asarray(hk), pk, th)       # complex-step friendly
    F = np.zeros((s,) + y.shape, dtype=dt)
    Y = np.zeros_like(F)
    for i in range(s):
        Yi = 1.0 * y
        for j in range(s):
            Yi = Yi + (hk * A[i, j]) * F[j]
        Y[i] = Yi
        F[i] = rhs.f(Yi, pk[i], th)
    return Y, F


def implicit_stages(rhs, A, y, hk, pk, th, tol=1e-7, max_iter=100):
    """ozone/integrators/time_marching.py:189-243.  Newton on
    R(Y) = Y - (A (x) I)(h F(Y)) - (U (x) I) y, initial value U y (:238), tolerance 1e-7 (:220).
    The residual is tested BEFORE each update (tutorial.ipynb:149: a linear problem reports
    "converged in 1 iterations").  All states are unknowns of one solver (:224-240).  Each
    trajectory of the batch is its own solve (converged ones are frozen).
    Returns Y, F (s, d, N), iteration counts (N,), converged flags (N,)."""
    s = A.shape[0]
    d, N = y.shape
    m = s * d
    Y = np.broadcast_to(y, (s, d, N)).copy()
    iters = np.zeros(N, dtype=np.int64)
    active = np.ones(N, dtype=bool)
    hk = np.broadcast_to(np.asarray(hk, dtype=float), (N,))
    F = np.zeros_like(Y)
    for it in range(max_iter + 1):
        for i in range(s):
            F[i] = rhs.f(Y[i], pk[i], th)
        R = Y - hk * np.einsum('ij,jdn->idn', A, F) - y
        rn = np.sqrt(np.sum(R.reshape(m, N) ** 2, axis=0))
        active = active & ~(rn < tol)
        if not active.any() or it == max_iter:
            break
        idx = np.nonzero(active)[0]
        # dR/dY: block (i, j) = delta_ij I - h a_ij J_f(Y_j)
        M = np.zeros((len(idx), s, d, s, d))
        for j in range(s):
            pj = pk[j] if pk[j].shape[-1] == 1 else pk[j][..., idx]
            tj = th if th.shape[-1] == 1 else th[..., idx]
            Jj = rhs.jac(Y[j][:, idx], pj, tj)          # (d, d, n_act)
            for i in range(s):
                M[:, i, :, j, :] -= (hk[idx] * A[i, j])[:, None, None] * np.moveaxis(Jj, -1, 0)
        M = M.reshape(len(idx), m, m) + np.eye(m)
        dY = np.linalg.solve(M, -np.moveaxis(R.reshape(m, N)[:, idx], -1, 0)[..., None])[..., 0]
        Y[:, :, idx] += np.moveaxis(dY, 0, -1).reshape(s, d, len(idx))
        iters[idx] += 1
    return Y, F, iters, ~active


def step(rhs, method, y, hk, pk, th, tol=1e-7, max_iter=100):
    """One GLM step: stages then ozone/integrators/time_marching.py:247-260
    y+ = h (B (x) I) F + (V (x) I) y  (V = 1)."""
    A, b, c, explicit = method
    if explicit:
        Y, F = explicit_stages(rhs, A, y, hk, pk, th)
        it = None
    else:
        Y, F, it, _ = implicit_stages(rhs, A, y, hk, pk, th, tol, max_iter)
    return y + hk * np.einsum('i,idn->dn', b, F), Y, F, it


def _as_method(method):
    return tableau(method) if isinstance(method, str) else method


def _norm_inputs(rhs, y0, h, p, th):
    def arr(x):
        x = np.asarray(x)
        return x if np.iscomplexobj(x) else x.astype(float)
    y0 = arr(y0)
    if y0.ndim == 1:
        y0 = y0[:, None]
    h = arr(h)
    n = h.shape[0]
    if p is None:
        p = np.zeros((n + 1, 0, 1))
    p = arr(p)
    if p.ndim == 2:
        p = p[:, :, None]
    if th is None:
        th = np.zeros((0, 1))
    th = arr(th)
    if th.ndim == 1:
        th = th[:, None]
    assert p.shape[0] == n + 1, 'dynamic parameters need one row per time point (collections.py:251-254)'
    return y0, h, p, th, n


def time_marching(rhs, method, y0, h, p=None, th=None, tol=1e-7, max_iter=100):
    """ozone/integrators/time_marching.py:6-59 + :97-141.
    history = [y_0; y_1..y_n]; final = y_n; profile[k] = g(y_k, p[k]); field = sum_{k=0..n} phi(y_k, p[k])
    with the RAW grid-point parameters (:30-32, :107)."""
    method = _as_method(method)
    y0, h, p, th, n = _norm_inputs(rhs, y0, h, p, th)
    ps = interpolate_params(p, method[2])
    y = y0.copy()
    hist = [y0.copy()]
    prof0, field = rhs.out(y0, p[0], th)
    prof = [prof0]
    field = field.copy()
    iters = []
    for k in range(n):
        y, _, _, it = step(rhs, method, y, _hk(h, k), ps[k], th, tol, max_iter)
        hist.append(y)
        pr, fi = rhs.out(y, p[k + 1], th)
        prof.append(pr)
        field = field + fi
        iters.append(it)
    N = y0.shape[1]
    prof = np.stack([np.broadcast_to(q, q.shape[:1] + (N,)) for q in prof])
    field = np.broadcast_to(field, field.shape[:1] + (N,)).copy()
    return dict(history=np.stack(hist), final=y, profile=prof, field=field,
                newton_iters=None if method[3] else np.stack(iters))


def checkpoint_layout(n, num_checkpoints=None):
    """ozone/integrators/time_marching_uniform.py:18-27: nc defaults to int(n**0.5); n % nc == 0
    is required; segment j covers steps [j*n/nc, (j+1)*n/nc)."""
    nc = int(n ** 0.5) if num_checkpoints is None else int(num_checkpoints)
    if n % nc != 0:
        raise ValueError(f'Number of timesteps ({n}) must be divisible by number of checkpoints {nc}')
    return nc, n // nc


def time_marching_checkpoints(rhs, method, y0, h, p=None, th=None, num_checkpoints=None, tol=1e-7,
                              max_iter=100):
    """ozone/integrators/time_marching_uniform.py:7-150: same arithmetic as time_marching run as
    nc outer x n_inner inner steps; only the outer-boundary states are retained for the reverse
    sweep (returned as 'checkpoints', (nc+1, d, N))."""
    method = _as_method(method)
    y0, h, p, th, n = _norm_inputs(rhs, y0, h, p, th)
    nc, ni = checkpoint_layout(n, num_checkpoints)
    ps = interpolate_params(p, method[2])
    y = y0.copy()
    cps = [y0.copy()]
    hist = [y0.copy()]
    prof0, field = rhs.out(y0, p[0], th)
    prof = [prof0]
    for o in range(nc):
        lo = o * ni
        for q in range(ni):
            k = lo + q
            y = step(rhs, method, y, _hk(h, k), ps[k], th, tol, max_iter)[0]
            hist.append(y)
            pr, fi = rhs.out(y, p[k + 1], th)
            prof.append(pr)
            field = field + fi
        cps.append(y)
    N = y0.shape[1]
    prof = np.stack([np.broadcast_to(q, q.shape[:1] + (N,)) for q in prof])
    field = np.broadcast_to(field, field.shape[:1] + (N,)).copy()
    return dict(history=np.stack(hist), final=y, profile=prof, field=field,
                checkpoints=np.stack(cps), num_checkpoints=nc, num_inner=ni)


# --------------------------------------------------------------------------------------------
def picard(rhs, method, y0, h, p=None, th=None, tol=1e-10, max_iter=2000, guess=None,
           state_blocks=None):
    """ozone/integrators/picard_iteration.py:9-65 (+ vector_based.py:9-53 for the outputs).

    Unknown Ybar (n, s, d, N).  hF = h_rep * f(Ybar, p_stage)               (:43-44)
    S = (I - shift (x) V)^-1 (B hF + y0vec) = running sum over time of [y0; B hF_0; ...]  (:17-19, :46)
    stages = A hF + U S                                                         (:46)
    residual = Ybar - stages;  update Ybar <- Ybar - residual                   (:51-52)
    States are relaxed one after another, each re-evaluating f (block Gauss-Seidel, :35);
    ``state_blocks`` lists the packed-component slices of the states in insertion order
    (default: one block = all components = Jacobi over components).  max_iter = 2000 (:35).
    Initial Ybar: user guess repeated per stage, else the broadcast IC (collections.py:125-133).
    History = S (:65).  Profile/field outputs are evaluated on the solved history with the raw
    parameters and the field is summed over time (vector_based.py:33-49)."""
    A, b, c, _ = _as_method(method)
    y0, h, p, th, n = _norm_inputs(rhs, y0, h, p, th)
    d, N = y0.shape
    s = A.shape[0]
    ps = interpolate_params(p, c)                               # (n, s, P, N|1)
    hh = (h if h.ndim == 2 else h[:, None]).reshape(n, 1, 1, -1)
    if guess is None:
        Yb = np.broadcast_to(y0, (n, s, d, N)).copy()
    else:
        Yb = np.repeat(np.asarray(guess, dtype=float).reshape(n, 1, d, -1), s, axis=1)
        Yb = np.broadcast_to(Yb, (n, s, d, N)).copy()
    blocks = [slice(0, d)] if state_blocks is None else state_blocks
    active = np.ones(N, dtype=bool)
    sweeps = np.zeros(N, dtype=np.int64)

    def evalF(Y):
        F = np.empty_like(Y)
        for i in range(s):
            F[:, i] = np.moveaxis(rhs.f(np.moveaxis(Y[:, i], 0, -1), np.moveaxis(ps[:, i], 0, -1), th[..., None]), -1, 0)
        return F

    def stages_of(Y):
        hF = hh * evalF(Y)
        inc = np.einsum('i,kidn->kdn', b, hF)
        S = np.concatenate([y0[None], y0[None] + np.cumsum(inc, axis=0)])      # (nt, d, N)
        return np.einsum('ij,kjdn->kidn', A, hF) + S[:-1, None], S

    S = None
    for it in range(max_iter + 1):
        rn = np.zeros(N)
        Ynew = Yb.copy()
        for blk in blocks:
            st, _ = stages_of(Ynew)
            res = Ynew[:, :, blk] - st[:, :, blk]
            rn = np.maximum(rn, np.sqrt(np.sum(res ** 2, axis=(0, 1, 2))))
            Ynew[:, :, blk] = st[:, :, blk]
        active = active & ~(rn < tol)
        if not active.any() or it == max_iter:
            break
        Yb[..., active] = Ynew[..., active]
        sweeps[active] += 1
    _, S = stages_of(Yb)
    pr, fi = rhs.out(np.moveaxis(S, 0, -1), np.moveaxis(p, 0, -1), th[..., None])
    prof = np.broadcast_to(np.moveaxis(pr, -1, 0), (n + 1, pr.shape[0], N))
    field = np.broadcast_to(np.broadcast_to(fi, fi.shape[:1] + (N, n + 1)).sum(axis=-1), (fi.shape[0], N))
    return dict(history=S, final=S[-1], profile=prof.copy(), field=field.copy(), stages=Yb,
                sweeps=sweeps, converged=~active)


# --------------------------------------------------------------------------------------------
def collocation_guess(y0, n, s, guess=None):
    """ozone/collections.py:125-133: user guess repeated per stage, else the broadcast IC."""
    d, N = y0.shape
    if guess is None:
        return np.broadcast_to(y0, (n, s, d, N)).copy(), np.broadcast_to(y0, (n, d, N)).copy()
    g = np.broadcast_to(np.asarray(guess, dtype=float).reshape(n, d, -1), (n, d, N))
    return np.repeat(g[:, None], s, axis=1).copy(), g.copy()


def collocation_scalers(stage_guess, state_guess, scaler=None):
    """ozone/collections.py:204-209."""
    if scaler is not None:
        return scaler, scaler
    st = 1 / np.average(np.abs(state_guess)) if np.abs(np.average(state_guess)) > 1e-7 else 1.0
    sg = 1 / np.average(np.abs(stage_guess)) if np.abs(np.average(stage_guess)) > 1e-7 else 1.0
    return sg, st


def collocation_residual(rhs, method, y0, h, Ydv, ydv, p=None, th=None):
    """ozone/integrators/collocation.py:33-50.  Ydv (n, s, d, N) stage unknowns, ydv (n, d, N)
    state unknowns y_1..y_n;  yfull = [y0; ydv] (collections.py:228-232).
    c1 = A hF + U yfull - Ydv            (n, s, d, N)
    c2 = (B hF + V yfull - yfull)[d:]    (n, d, N)   i.e. h sum_i b_i F_ki + y_k - y_{k+1}."""
    A, b, c, _ = _as_method(method)
    y0, h, p, th, n = _norm_inputs(rhs, y0, h, p, th)
    s = A.shape[0]
    ps = interpolate_params(p, c)
    hh = (h if h.ndim == 2 else h[:, None]).reshape(n, 1, 1, -1)
    F = np.empty_like(Ydv)
    for i in range(s):
        F[:, i] = np.moveaxis(rhs.f(np.moveaxis(Ydv[:, i], 0, -1), np.moveaxis(ps[:, i], 0, -1), th[..., None]), -1, 0)
    hF = hh * F
    yfull = np.concatenate([np.broadcast_to(y0, ydv.shape[1:])[None], ydv])
    c1 = np.einsum('ij,kjdn->kidn', A, hF) + yfull[:-1, None] - Ydv
    c2 = np.einsum('i,kidn->kdn', b, hF) + yfull[:-1] - yfull[1:]
    return c1, c2
