"""Oracle copy of the Butcher tableaux (TEST INFRASTRUCTURE, see oracle/__init__.py).

Follows ``ozone/glms/GLMs.py:45-68`` (registry) and the coefficient files under
``ozone/glms/methods/runge_kutta/``.  Kept separate from ``ozone_b200/glms.py`` on purpose:
``tests/test_tableaux.py`` compares the two and re-derives the collocation families
(Gauss, Radau IIA, Lobatto IIIA) from their abscissae as a third opinion.
"""
import numpy as np
from fractions import Fraction as Fr

sqrt = np.sqrt


def _f(x):
    return float(x)


def tableau(name):
    """Return (A (s,s), b (s,), c (s,), explicit) for an ``ozone.methods`` name."""
    n = {'AB1': 'ForwardEuler', 'AM1': 'BackwardEuler', 'BDF1': 'BackwardEuler'}.get(name, name)
    if n == 'ForwardEuler':                       # explicit_runge_kutta.py:9-14
        A, b = [[0.]], [1.]
    elif n == 'ExplicitMidpoint':                 # :17-40
        A, b = [[0, 0], [_f(Fr(1, 2)), 0]], [0, 1]
    elif n == 'HeunsMethod':                      # :61-74
        A, b = [[0, 0], [1, 0]], [.5, .5]
    elif n == 'RalstonsMethod':                   # :77-90
        A, b = [[0, 0], [2 / 3, 0]], [1 / 4, 3 / 4]
    elif n == 'KuttaThirdOrder':                  # :93-117
        A, b = [[0, 0, 0], [.5, 0, 0], [-1, 2, 0]], [1 / 6, 4 / 6, 1 / 6]
    elif n == 'RK4':                              # :138-163
        A = np.zeros((4, 4)); A[1, 0] = .5; A[2, 1] = .5; A[3, 2] = 1.
        b = [1 / 6, 1 / 3, 1 / 3, 1 / 6]
    elif n == 'RK6':                              # :184-215, s = 1
        r = sqrt(5)
        A = np.zeros((7, 7))
        A[1, :1] = (5 - r) / 10
        A[2, :2] = [-r / 10, (5 + 2 * r) / 10]
        A[3, :3] = [(-15 + 7 * r) / 20, (-1 + r) / 4, (15 - 7 * r) / 10]
        A[4, [0, 2, 3]] = [(5 - r) / 60, 1 / 6, (15 + 7 * r) / 60]
        A[5, [0, 2, 3, 4]] = [(5 + r) / 60, (9 - 5 * r) / 12, 1 / 6, (-5 + 3 * r) / 10]
        A[6, [0, 2, 3, 4, 5]] = [1 / 6, (-55 + 25 * r) / 12, (-25 - 7 * r) / 12, 5 - 2 * r, (5 + r) / 2]
        b = np.zeros(7); b[[0, 4, 5, 6]] = [1 / 12, 5 / 12, 5 / 12, 1 / 12]
    elif n == 'BackwardEuler':                    # implicit_runge_kutta.py:8-13
        A, b = [[1.]], [1.]
    elif n in ('ImplicitMidpoint', 'GaussLegendre2'):   # :16-21 ; gauss_legendre.py:9-12
        A, b = [[.5]], [1.]
    elif n in ('Trapezoidal', 'Lobatto2'):        # :24-30 ; lobatto.py:9-13
        A, b = [[0, 0], [.5, .5]], [.5, .5]
    elif n == 'GaussLegendre4':                   # gauss_legendre.py:14-18
        q = sqrt(3) / 6
        A, b = [[.25, .25 - q], [.25 + q, .25]], [.5, .5]
    elif n == 'GaussLegendre6':                   # :19-24
        q = sqrt(15)
        A = [[5 / 36, 2 / 9 - q / 15, 5 / 36 - q / 30],
             [5 / 36 + q / 24, 2 / 9, 5 / 36 - q / 24],
             [5 / 36 + q / 30, 2 / 9 + q / 15, 5 / 36]]
        b = [5 / 18, 4 / 9, 5 / 18]
    elif n == 'Lobatto4':                         # lobatto.py:14-19
        A, b = [[0, 0, 0], [5 / 24, 1 / 3, -1 / 24], [1 / 6, 2 / 3, 1 / 6]], [1 / 6, 2 / 3, 1 / 6]
    elif n == 'RadauI3':                          # radau.py:8-12
        A, b = [[1 / 4, -1 / 4], [1 / 4, 5 / 12]], [1 / 4, 3 / 4]
    elif n == 'RadauI5':                          # :13-18
        q = sqrt(6)
        A = [[1 / 9, (-1 - q) / 18, (-1 + q) / 18],
             [1 / 9, 11 / 45 + 7 * q / 360, 11 / 45 - 43 * q / 360],
             [1 / 9, 11 / 45 + 43 * q / 360, 11 / 45 - 7 * q / 360]]
        b = [1 / 9, 4 / 9 + q / 36, 4 / 9 - q / 36]
    elif n == 'RadauII3':                         # :22-26
        A, b = [[5 / 12, -1 / 12], [3 / 4, 1 / 4]], [3 / 4, 1 / 4]
    elif n == 'RadauII5':                         # :27-32
        q = sqrt(6)
        A = [[11 / 45 - 7 * q / 360, 37 / 225 - 169 * q / 1800, -2 / 225 + q / 75],
             [37 / 225 + 169 * q / 1800, 11 / 45 + 7 * q / 360, -2 / 225 - q / 75],
             [4 / 9 - q / 36, 4 / 9 + q / 36, 1 / 9]]
        b = [4 / 9 - q / 36, 4 / 9 + q / 36, 1 / 9]
    else:
        raise KeyError(name)
    A = np.atleast_2d(np.asarray(A, dtype=float))
    b = np.asarray(b, dtype=float).reshape(-1)
    c = A.dot(np.ones((A.shape[0], 1))).reshape(-1)   # GLMs.py:29  C = A.dot(ones): the same operation, so the same bits
    explicit = bool(np.linalg.norm(np.tril(A, -1) - A) < 1e-15)   # GLMs.py:37-39
    return A, b, c, explicit


ALL_NAMES = ['ForwardEuler', 'BackwardEuler', 'ExplicitMidpoint', 'ImplicitMidpoint', 'KuttaThirdOrder',
             'RK4', 'RK6', 'RalstonsMethod', 'HeunsMethod', 'GaussLegendre2', 'GaussLegendre4',
             'GaussLegendre6', 'Lobatto2', 'Lobatto4', 'RadauI3', 'RadauI5', 'RadauII3', 'RadauII5',
             'Trapezoidal', 'AB1', 'AM1', 'BDF1']
