"""Shared seeded inputs for the parity tests: the reference's own test problems
(ozone/tests/test_correctness/*/utils.py) and the BASELINE.json configs at oracle-friendly sizes.
All arrays follow the oracle conventions: y0 (d, N), h (n,), p (nt, P, N|1), th (S, N|1)."""
import numpy as np


def test1_inputs():
    """ozone/tests/test_correctness/test_1_simple/utils.py:9-28: y' = -y, y0 = .5, 29 x 0.01."""
    return dict(rhs='affine', y0=np.array([[0.5]]), h=np.full(29, 0.01), p=None, th=np.array([[-1.0], [0.0]]))


def tutorial_inputs():
    """examples/simple_examples/tutorial.ipynb: y' = .5 y - 3, y0 = 1, 10 x 0.1."""
    return dict(rhs='affine', y0=np.array([[1.0]]), h=np.full(10, 0.1), p=None, th=np.array([[0.5], [-3.0]]))


def test2_inputs():
    """ozone/tests/test_correctness/test_2_general/utils.py:11-41 (nt = 31, h = 1e-3)."""
    nt = 31
    t = np.arange(nt)
    a, b, g = 1.0 + t / nt / 5, 0.5 + t / nt / 5, 2.0 + t / nt / 5
    e = np.array([[0.3, 0.2], [0.1, -2.6]])[None] + (t / nt / 5)[:, None, None]
    co = np.ones(nt) / nt
    p = np.concatenate([a[:, None], b[:, None], g[:, None], e.reshape(nt, 4), co[:, None]], axis=1)[:, :, None]
    y0 = np.array([2.0, 2.0, 2.0, 1.0, -1.0, -3.0])[:, None]
    return dict(rhs='general3', y0=y0, h=np.full(nt - 1, 1e-3), p=p, th=np.array([[0.5]]))


def test3_inputs(nt=30, N=500):
    """ozone/tests/test_correctness/test_3_large/utils.py:6-38."""
    return dict(rhs='quad_decay', y0=0.5 * np.ones((1, N)), h=np.full(nt - 1, 0.01),
                p=np.linspace(0.5, 0.8, nt).reshape(nt, 1, 1), th=np.linspace(0.1, 1.1, N)[None])


def lotka_volterra_inputs(N=1, nt=101, seed=0, shared_a=True):
    """BASELINE config C1 (N = 1) / C4 (batched): examples/simple_examples/lotka_volterra.py with
    num_times = 101, h = 0.009 (same horizon T = 0.9)."""
    rng = np.random.default_rng(seed)
    t = np.arange(nt)
    a = (1.0 + t / nt / 5.0).reshape(nt, 1, 1)
    if N == 1:
        y0 = np.array([[2.0], [2.0]])
        c = np.array([[0.5]])
    else:
        y0 = rng.uniform(1.5, 2.5, size=(2, N))
        c = rng.uniform(0.4, 0.6, size=(1, N))
    if not shared_a:
        a = a * (1.0 + 0.05 * rng.standard_normal((1, 1, N)))
    return dict(rhs='lotka_volterra', y0=y0, h=np.full(nt - 1, 0.9 / (nt - 1)), p=a, th=c)


def van_der_pol_inputs(N=64, nt=40, seed=0):
    """BASELINE config C2: ozone/paper_examples/van_der_pol (nt = 40, tf = 15, IC (1,1,0),
    u in [-0.75, 1]) with one control sequence per trajectory."""
    rng = np.random.default_rng(seed)
    u = rng.uniform(-0.75, 1.0, size=(nt, 1, N))
    y0 = np.tile(np.array([[1.0], [1.0], [0.0]]), (1, N))
    return dict(rhs='van_der_pol', y0=y0, h=np.full(nt - 1, 15.0 / (nt - 1)), p=u, th=None)


ASCENT_IC = np.array([0.061663494666509, 0.055249439788994, -1.004110522066614,
                      -0.465976472038572, -0.416909510597800, -0.084973690970470,
                      11846.599614336959 / 15103.])   # ascent_system/main_model.py:72-79


def ascent_inputs(N=16, nt=101, seed=0, tf=0.1):
    """BASELINE config C5: ozone/paper_examples/ascent_system/main_model.py:63-79."""
    rng = np.random.default_rng(seed)
    u = rng.uniform(-1.0, 1.0, size=(nt, 3, N))
    y0 = np.tile(ASCENT_IC[:, None], (1, N))
    return dict(rhs='lunar_ascent', y0=y0, h=np.full(nt - 1, tf / (nt - 1)), p=u, th=None)


def pde_inputs(nx=200, n=16, seed=0):
    """BASELINE config C3 scaled down: pde_control.py:99-108 with dt = 0.4 dx^2."""
    rng = np.random.default_rng(seed)
    dx = np.pi / nx
    u = rng.uniform(-0.6, 0.0, size=(n + 1, 1, 1))
    return dict(rhs='diffusion_reaction', y0=0.5 * np.ones((nx, 1)), h=np.full(n, 0.4 * dx * dx), p=u, th=None,
                nx=nx)
