"""Argument checks (TypeError naming the offending argument, as ``ozone/utils/general.py`` does) and the
conversion of user values to the engine's variable type."""
import numpy as np
import torch

_DEFAULT = {'device': None, 'dtype': torch.float64}


def default_device():
    if _DEFAULT['device'] is None:
        _DEFAULT['device'] = torch.device('cuda', torch.cuda.current_device()) if torch.cuda.is_available() \
            else torch.device('cpu')
    return _DEFAULT['device']


def set_default_dtype(dtype):
    """fp64 (default) or fp32: dtype given to values that are not already tensors."""
    if dtype not in (torch.float64, torch.float32):
        raise TypeError('dtype must be torch.float64 or torch.float32')
    _DEFAULT['dtype'] = dtype


def get_type_string(value):
    return repr(type(value).__name__)


def _type_check(expected_label, accept):
    """Build a checker that raises TypeError naming the argument (the reference's convention, general.py:5-26)."""
    def check(value, argname):
        if not accept(value):
            raise TypeError(f"argument '{argname}' must be {expected_label}, got {get_type_string(value)}")
    return check


def _is_method(v):
    from .methods import _Method
    return isinstance(v, _Method)


check_if_string = _type_check('a string', lambda v: isinstance(v, str))
check_if_bool = _type_check('a bool', lambda v: isinstance(v, bool))
check_if_int = _type_check('an int', lambda v: isinstance(v, (int, np.int32, np.int64)) and not isinstance(v, bool))
check_if_method = _type_check('an instance of ozone.methods._Method', _is_method)


def variablize(value, argname='value'):
    """The reference turns numbers / lists / arrays into ``csdl.Variable`` and rejects anything else
    with TypeError (``ozone/utils/general.py:34-39``).  Here the variable type is ``torch.Tensor``:
    tensors pass through (keeping their autograd history), numbers / lists / arrays become tensors
    of the default dtype on the default device.  A python scalar becomes shape (1,), as a scalar
    csdl.Variable does."""
    if isinstance(value, torch.Tensor):
        return value
    if isinstance(value, list):
        value = np.array(value)
    if isinstance(value, (bool, str)) or not isinstance(value, (int, float, np.ndarray, np.floating, np.integer)):
        raise TypeError(f'expected a tensor, array or number for argument \'{argname}\' but got {get_type_string(value)}')
    arr = np.asarray(value, dtype=np.float64)
    if arr.ndim == 0:
        arr = arr.reshape(1)
    return torch.tensor(arr, dtype=_DEFAULT['dtype'], device=default_device())
