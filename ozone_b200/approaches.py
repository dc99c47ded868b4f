"""Solution approaches (``ozone/approaches.py:9-30``): same class names, same ``.name`` strings
(they are used as file names by the reference's harness)."""
from typing import Optional


class _Approach(object):
    name: str = 'None'
    kind: str = 'none'

    def __init__(self) -> None:
        pass


class TimeMarching(_Approach):
    name = 'TimeMarching'
    kind = 'time_marching'


class TimeMarchingCheckpoints(_Approach):
    name = 'TimeMarching (Checkpointing)'
    kind = 'time_marching_checkpoints'

    def __init__(self, num_checkpoints: Optional[int] = None) -> None:
        super().__init__()
        self.num_checkpoints = num_checkpoints


class PicardIteration(_Approach):
    name = 'Picard Iteration'
    kind = 'picard'

    def __init__(self, tolerance: Optional[float] = None, max_iter: int = 2000, relaxation: str = 'gauss_seidel') -> None:
        # max_iter: ozone/integrators/picard_iteration.py:35.  The tolerance is a csdl default the
        # reference does not set; here it defaults to 1e-10 (fp64) / 1e-4 (fp32).
        # relaxation: 'gauss_seidel' relaxes the states one after another, re-evaluating f for each (what csdl's
        # GaussSeidel solver does, :35); 'jacobi' relaxes all states from one evaluation of f -- same fixed point,
        # one RHS pass per sweep instead of one per state.
        super().__init__()
        if relaxation not in ('gauss_seidel', 'jacobi'):
            raise ValueError("relaxation must be 'gauss_seidel' or 'jacobi'")
        self.tolerance, self.max_iter, self.relaxation = tolerance, max_iter, relaxation


class Collocation(_Approach):
    name = 'Collocation'
    kind = 'collocation'
