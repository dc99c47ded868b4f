"""Time spans (``ozone/timespan.py``).  ``finalize() -> (time_vector (None), step_vector, num_times)``."""
import torch

from .utils import variablize, check_if_int


def check_if_scalar(value, argname):
    """one element of any shape -> shape (1,); anything else is a ValueError"""
    if value.numel() == 1:
        return value.reshape((1,))
    raise ValueError(f"argument '{argname}' must hold exactly one value, its shape is {tuple(value.shape)}")


def check_if_vector(value, argname):
    if value.dim() != 1:
        raise ValueError(f"argument '{argname}' must be one-dimensional, its shape is {tuple(value.shape)}")
    return value


class TimeSpan(object):
    def __init__(self) -> None:
        self.step_vector = None
        self.time_vector = None

    def finalize(self):
        """ozone/timespan.py:25-33."""
        steps = self.step_vector
        if steps.shape[0] < 1:
            raise ValueError('a time span needs at least one step (two time points)')
        return self.time_vector, steps, steps.shape[0] + 1


class StepVector(TimeSpan):
    """``StepVector(start, step_vector)`` (ozone/timespan.py:63-76).  ``step_vector`` is 1-D (n,)
    -- shared by every trajectory of a batch -- or, as an extension, (n, N) with one column per
    trajectory."""

    def __init__(self, start, step_vector) -> None:
        super().__init__()
        start = check_if_scalar(variablize(start, 'start'), 'start')
        step_vector = variablize(step_vector, 'step_vector')
        if step_vector.dim() != 2:
            step_vector = check_if_vector(step_vector, 'step_vector')
        self.start = start
        self.step_vector = step_vector


class StartEnd(TimeSpan):
    """``StartEnd(start, end, num_steps)``: ``num_steps`` equal steps of (end - start) / num_steps.

    DEVIATION: the reference builds ``linspace(start, end, num_steps) - start``
    (ozone/timespan.py:60), i.e. cumulative offsets rather than step sizes; it is untested
    (ozone/tests/test_api_timespans.py is empty) and unused.  The evident intent is implemented."""

    def __init__(self, start, end, num_steps: int) -> None:
        super().__init__()
        start = check_if_scalar(variablize(start, 'start'), 'start')
        end = check_if_scalar(variablize(end, 'end'), 'end')
        check_if_int(num_steps, 'num_steps')
        self.start = start
        self.step_vector = ((end - start) / num_steps).expand(num_steps)
