"""Runge-Kutta tableaux as data, keyed by the ``ozone.methods`` names.

Reference: the registry ``ozone/glms/GLMs.py:45-68`` (name -> tableau) and the coefficient
sources ``ozone/glms/methods/runge_kutta/explicit_runge_kutta.py``, ``implicit_runge_kutta.py``,
``gauss_legendre.py``, ``lobatto.py``, ``radau.py``.  Every registered method is a one-value
Runge-Kutta scheme: ``U = ones(s,1)``, ``V = [[1]]`` (``runge_kutta.py:7-19``), so the device
kernels only ever see ``(A, B, C=A.1, explicit)``.

A tableau is "explicit" when ``A`` is strictly lower triangular within 1e-15
(``GLMs.py:37-39``); anything else (including Trapezoidal / Lobatto, whose first row is zero)
takes the implicit (Newton) path.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

_S3, _S5, _S6, _S15 = math.sqrt(3.0), math.sqrt(5.0), math.sqrt(6.0), math.sqrt(15.0)


def _rk6_rows():
    # explicit_runge_kutta.py:184-205 with s = +1
    r = _S5
    A = np.zeros((7, 7))
    A[1, 0] = (5 - r) / 10
    A[2, 0], A[2, 1] = -r / 10, (5 + 2 * r) / 10
    A[3, 0], A[3, 1], A[3, 2] = (-15 + 7 * r) / 20, (-1 + r) / 4, (15 - 7 * r) / 10
    A[4, 0], A[4, 2], A[4, 3] = (5 - r) / 60, 1 / 6, (15 + 7 * r) / 60
    A[5, 0], A[5, 2], A[5, 3], A[5, 4] = (5 + r) / 60, (9 - 5 * r) / 12, 1 / 6, (-5 + 3 * r) / 10
    A[6, 0] = 1 / 6
    A[6, 2], A[6, 3], A[6, 4], A[6, 5] = (-55 + 25 * r) / 12, (-25 - 7 * r) / 12, 5 - 2 * r, (5 + r) / 2
    b = [1 / 12, 0, 0, 0, 5 / 12, 5 / 12, 1 / 12]
    return A.tolist(), b


_RK6_A, _RK6_B = _rk6_rows()

# name -> (A rows, b weights)
_TABLEAUX = {
    'ForwardEuler': ([[0.0]], [1.0]),
    'ExplicitMidpoint': ([[0, 0], [0.5, 0]], [0, 1]),
    'HeunsMethod': ([[0, 0], [1, 0]], [0.5, 0.5]),
    'RalstonsMethod': ([[0, 0], [2 / 3, 0]], [0.25, 0.75]),
    'KuttaThirdOrder': ([[0, 0, 0], [0.5, 0, 0], [-1, 2, 0]], [1 / 6, 4 / 6, 1 / 6]),
    'RK4': ([[0, 0, 0, 0], [0.5, 0, 0, 0], [0, 0.5, 0, 0], [0, 0, 1, 0]], [1 / 6, 1 / 3, 1 / 3, 1 / 6]),
    'RK6': (_RK6_A, _RK6_B),
    'BackwardEuler': ([[1.0]], [1.0]),
    'ImplicitMidpoint': ([[0.5]], [1.0]),
    'Trapezoidal': ([[0, 0], [0.5, 0.5]], [0.5, 0.5]),
    'GaussLegendre2': ([[0.5]], [1.0]),
    'GaussLegendre4': ([[1 / 4, 1 / 4 - _S3 / 6], [1 / 4 + _S3 / 6, 1 / 4]], [0.5, 0.5]),
    'GaussLegendre6': (
        [[5 / 36, 2 / 9 - _S15 / 15, 5 / 36 - _S15 / 30],
         [5 / 36 + _S15 / 24, 2 / 9, 5 / 36 - _S15 / 24],
         [5 / 36 + _S15 / 30, 2 / 9 + _S15 / 15, 5 / 36]],
        [5 / 18, 4 / 9, 5 / 18]),
    'Lobatto2': ([[0, 0], [0.5, 0.5]], [0.5, 0.5]),
    'Lobatto4': ([[0, 0, 0], [5 / 24, 1 / 3, -1 / 24], [1 / 6, 2 / 3, 1 / 6]], [1 / 6, 2 / 3, 1 / 6]),
    'RadauI3': ([[1 / 4, -1 / 4], [1 / 4, 5 / 12]], [1 / 4, 3 / 4]),
    'RadauI5': (
        [[1 / 9, (-1 - _S6) / 18, (-1 + _S6) / 18],
         [1 / 9, 11 / 45 + 7 * _S6 / 360, 11 / 45 - 43 * _S6 / 360],
         [1 / 9, 11 / 45 + 43 * _S6 / 360, 11 / 45 - 7 * _S6 / 360]],
        [1 / 9, 4 / 9 + _S6 / 36, 4 / 9 - _S6 / 36]),
    'RadauII3': ([[5 / 12, -1 / 12], [3 / 4, 1 / 4]], [3 / 4, 1 / 4]),
    'RadauII5': (
        [[11 / 45 - 7 * _S6 / 360, 37 / 225 - 169 * _S6 / 1800, -2 / 225 + _S6 / 75],
         [37 / 225 + 169 * _S6 / 1800, 11 / 45 + 7 * _S6 / 360, -2 / 225 - _S6 / 75],
         [4 / 9 - _S6 / 36, 4 / 9 + _S6 / 36, 1 / 9]],
        [4 / 9 - _S6 / 36, 4 / 9 + _S6 / 36, 1 / 9]),
}
# aliases (GLMs.py:65-67)
_TABLEAUX['AB1'] = _TABLEAUX['ForwardEuler']
_TABLEAUX['AM1'] = _TABLEAUX['BackwardEuler']
_TABLEAUX['BDF1'] = _TABLEAUX['BackwardEuler']

MAX_STAGES = 7  # RK6; also the size of the device-side tableau storage


@dataclass(frozen=True)
class GLMMethod:
    """Same fields as the reference's ``GLMMethod`` (``ozone/glms/GLMs.py:19-28``)."""
    A: np.ndarray          # (s, s)
    B: np.ndarray          # (1, s)
    U: np.ndarray          # (s, 1) ones
    V: np.ndarray          # (1, 1) one
    explicit: bool
    num_stages: int
    C: np.ndarray          # (s, 1) = A @ 1


def get_integration_method(name: str) -> GLMMethod:
    """``ozone/glms/GLMs.py:30-42``: look the tableau up by name and classify it."""
    rows, weights = _TABLEAUX[name]
    A = np.atleast_2d(np.asarray(rows, dtype=np.float64))
    B = np.atleast_2d(np.asarray(weights, dtype=np.float64))
    s = A.shape[0]
    if A.shape != (s, s) or B.shape != (1, s):
        raise ValueError(f'inconsistent tableau {name}: A {A.shape}, B {B.shape}')
    explicit = bool(np.linalg.norm(np.tril(A, -1) - A) < 1e-15)
    return GLMMethod(A=A, B=B, U=np.ones((s, 1)), V=np.ones((1, 1)), explicit=explicit,
                     num_stages=s, C=A @ np.ones((s, 1)))


method_names = tuple(_TABLEAUX)
