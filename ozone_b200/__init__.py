"""ozone_b200 -- B200-native drop-in for the solve-and-differentiate path of LSDOlab/ozone.

Same surface as ``import ozone`` (``ozone/__init__.py``): ``ODEProblem``, ``methods``,
``approaches``, ``timespans`` (alias of the ``timespan`` module, both spellings are used by the
reference's callers), ``ODEVars``, ``IntegratorOutputs``; plus ``rhs`` (registered device
right-hand sides).  Values are ``torch.Tensor`` on a CUDA device; derivatives come from
``torch.autograd`` through the adjoint kernels.  There is no CPU fallback.
"""
__version__ = '0.1.0'

from . import _native          # loads liboz_b200.so; raises if it has not been built
from .odeproblem import ODEProblem
from . import timespan
from . import timespan as timespans
from . import approaches
from . import methods
from . import rhs
from . import ops
from .func_wrapper import ODEVars
from .collections import IntegratorOutputs
from .utils import set_default_dtype
