"""Argument checks with the reference's error types (``ozone/utils/general.py:2-39``)."""
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
    return f'\'{type(value).__name__}\''


def check_if_string(value, argname):
    if not isinstance(value, str):
        raise TypeError(f'expected type \'string\' for argument \'{argname}\' but got {get_type_string(value)}')


def check_if_method(value, argname):
    from .methods import _Method
    if not isinstance(value, _Method):
        raise TypeError(f'expected type \'ozone.methods._Method\' for argument \'{argname}\' but got {get_type_string(value)}')


def check_if_bool(value, argname):
    if not isinstance(value, bool):
        raise TypeError(f'expected type \'bool\' for argument \'{argname}\' but got {get_type_string(value)}')


def check_if_int(value, argname):
    if isinstance(value, bool) or not isinstance(value, (int, np.int32, np.int64)):
        raise TypeError(f'expected type \'int\' for argument \'{argname}\' but got {get_type_string(value)}')


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
