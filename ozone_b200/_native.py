"""ctypes binding of ``liboz_b200.so`` (the C ABI declared in ``include/ozone_b200.h``).

There is deliberately no fallback: if the shared library is missing or fails to load, importing
this module raises.  Build it with ``python -c "import __graft_entry__ as g; g.build()"`` or
``make -C ozone_b200/csrc``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('OZONE_B200_LIB') or os.path.join(_HERE, 'liboz_b200.so')   # env override: A/B builds

OZ_MAX_STAGES = 7
OZ_F64, OZ_F32 = 0, 1
STATUS_NOT_CONVERGED, STATUS_SINGULAR, STATUS_NONFINITE = 1, 2, 4
GRID_SKIP_IC_FIELD, GRID_SKIP_TERMINAL_FIELD, GRID_NO_CONTROL_COST = 1, 2, 4

RHS_IDS = {
    'lotka_volterra': 0,
    'van_der_pol': 1,
    'lunar_ascent': 2,
    'affine': 3,
    'quad_decay': 4,
    'general3': 5,
    'diffusion_reaction': 6,
}


class OzProblem(C.Structure):
    """Field-for-field mirror of ``struct oz_problem``."""
    _fields_ = [
        ('rhs_id', C.c_int32), ('dtype', C.c_int32), ('num_stages', C.c_int32), ('is_explicit', C.c_int32),
        ('A', C.c_double * (OZ_MAX_STAGES * OZ_MAX_STAGES)), ('B', C.c_double * OZ_MAX_STAGES),
        ('C', C.c_double * OZ_MAX_STAGES),
        ('N', C.c_int64), ('n', C.c_int32), ('num_inner', C.c_int32),
        ('h_shared', C.c_int32), ('p_shared', C.c_int32), ('th_shared', C.c_int32), ('d_override', C.c_int32),
        ('newton_tol', C.c_double), ('newton_max_iter', C.c_int32), ('reserved0', C.c_int32),
        ('rhs_cfg', C.c_double * 8),
        ('num_state_blocks', C.c_int32), ('state_block_lo', C.c_int32 * 9),
        ('grid_x_off', C.c_int64), ('grid_nx_global', C.c_int64), ('grid_own_lo', C.c_int64), ('grid_own_hi', C.c_int64), ('grid_flags', C.c_int64),
        ('y0', C.c_void_p), ('h', C.c_void_p), ('p', C.c_void_p), ('th', C.c_void_p),
        ('hist', C.c_void_p), ('ckpt', C.c_void_p), ('yfinal', C.c_void_p), ('prof', C.c_void_p),
        ('field', C.c_void_p), ('status', C.c_void_p), ('iters', C.c_void_p),
        ('hist_bar', C.c_void_p), ('final_bar', C.c_void_p), ('prof_bar', C.c_void_p), ('field_bar', C.c_void_p),
        ('y0_bar', C.c_void_p), ('h_bar', C.c_void_p), ('p_bar', C.c_void_p), ('th_bar', C.c_void_p),
        ('ws', C.c_void_p), ('ws_bytes', C.c_size_t), ('stream', C.c_void_p), ('stages', C.c_void_p),
    ]


# every symbol include/ozone_b200.h declares: name -> (restype, argtypes)
_PB = C.POINTER(OzProblem)
_IP = C.POINTER(C.c_int)
SYMBOLS = {
    'oz_version': (C.c_int, []),
    'oz_last_error': (C.c_char_p, []),
    'oz_rhs_info': (C.c_int, [C.c_int, _IP, _IP, _IP, _IP, _IP, _IP]),
    'oz_register_rhs_plugin': (C.c_int, [C.c_char_p]),
    'oz_fma_probe': (C.c_double, [C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    'oz_rhs_eval': (C.c_int, [_PB, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'oz_rhs_out_vjp': (C.c_int, [_PB, C.c_int64] + [C.c_void_p] * 7),
    'oz_rhs_jac_eval': (C.c_int, [_PB, C.c_int64, C.c_int64, C.c_int32] + [C.c_void_p] * 5),
    'oz_rhs_jac_pattern': (C.c_int, [C.c_int, C.c_void_p, C.c_void_p]),
    'oz_tm_forward_workspace_bytes': (C.c_size_t, [_PB]),
    'oz_tm_forward': (C.c_int, [_PB]),
    'oz_tm_adjoint_workspace_bytes': (C.c_size_t, [_PB]),
    'oz_tm_adjoint': (C.c_int, [_PB]),
    'oz_static_scratch_bytes': (C.c_size_t, []),
    'oz_tm_adjoint_static': (C.c_int, [_PB]),
    'oz_peer_halo_push': (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_uint64, C.c_void_p]),
    'oz_peer_halo_wait': (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]),
    'oz_peer_halo_exchange': (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_int64] + [C.c_void_p] * 11 + [C.c_uint64, C.c_void_p, C.c_void_p]),
    'oz_peer_epoch_advance': (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p]),
    'oz_peer_sum': (C.c_int, [C.c_int, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                              C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
    'oz_picard_workspace_bytes': (C.c_size_t, [_PB]),
    'oz_picard_forward': (C.c_int, [_PB, C.c_void_p, C.c_double, C.c_int]),
    'oz_colloc_residual': (C.c_int, [_PB, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'oz_colloc_vjp_workspace_bytes': (C.c_size_t, [_PB]),
    'oz_colloc_vjp': (C.c_int, [_PB, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f'{LIB_PATH} is missing: the CUDA library has not been built '
            f'(run `make -C {os.path.join(_HERE, "csrc")}`); ozone_b200 has no CPU fallback')
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)          # AttributeError if the ABI and the header drifted apart
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


class OzError(RuntimeError):
    pass


def check(rc: int, what: str):
    if rc != 0:
        raise OzError(f'{what} failed (rc={rc}): {lib.oz_last_error().decode()}')


def rhs_info(rhs_id: int):
    vals = [C.c_int() for _ in range(6)]
    check(lib.oz_rhs_info(rhs_id, *[C.byref(v) for v in vals]), 'oz_rhs_info')
    d, P, S, nprof, nfield, kind = (v.value for v in vals)
    return dict(d=d, P=P, S=S, nprof=nprof, nfield=nfield, kind=kind)
