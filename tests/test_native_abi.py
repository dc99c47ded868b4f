"""The C-ABI library loads without a GPU, exports every symbol include/ozone_b200.h declares, and the
ctypes mirror of `struct oz_problem` has the layout the C compiler gives it.  No compute calls."""
import ctypes as C
import os
import re
import subprocess
import sys
import tempfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, 'include', 'ozone_b200.h')


def _declared_symbols():
    text = open(HEADER).read()
    return sorted(set(re.findall(r'OZ_API\s+[\w\s\*]+?\b(oz_\w+)\s*\(', text)))


def test_every_declared_symbol_is_exported_and_bound():
    from ozone_b200 import _native as nat
    declared = _declared_symbols()
    assert len(declared) >= 15
    assert sorted(nat.SYMBOLS) == declared, 'include/ozone_b200.h and ozone_b200/_native.py drifted apart'
    for name in declared:
        assert getattr(nat.lib, name) is not None
    out = subprocess.run(['nm', '-D', '--defined-only', nat.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r' T (oz_\w+)', out))
    assert set(declared) <= exported
    assert nat.lib.oz_version() >= 100


def test_struct_layout_matches_the_c_compiler():
    from ozone_b200 import _native as nat
    fields = [f[0] for f in nat.OzProblem._fields_]
    prog = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void) {',
            '  printf("%zu\\n", sizeof(oz_problem));']
    prog += [f'  printf("%zu\\n", offsetof(oz_problem, {f}));' for f in fields]
    prog += ['  return 0; }']
    with tempfile.TemporaryDirectory() as d:
        src, exe = os.path.join(d, 't.c'), os.path.join(d, 't')
        open(src, 'w').write('\n'.join(prog))
        subprocess.run(['gcc', '-std=c99', src, '-o', exe], check=True)
        nums = [int(x) for x in subprocess.run([exe], capture_output=True, text=True, check=True).stdout.split()]
    assert nums[0] == C.sizeof(nat.OzProblem)
    for f, off in zip(fields, nums[1:]):
        assert getattr(nat.OzProblem, f).offset == off, f


def test_rhs_registry_sizes_without_a_gpu():
    from ozone_b200 import _native as nat, rhs
    import math
    for handle in (rhs.lotka_volterra, rhs.van_der_pol, rhs.lunar_ascent, rhs.affine, rhs.quad_decay, rhs.general3):
        info = nat.rhs_info(handle.rhs_id)
        size = lambda decl: sum(int(math.prod(s)) if s else 1 for _, s in decl)
        assert info['d'] == size(handle.states) and info['P'] == size(handle.dynamic_parameters)
        assert info['S'] == size(handle.static_parameters)
        assert info['nprof'] == size(handle.profile_outputs) and info['nfield'] == size(handle.field_outputs)
        assert info['kind'] == 0
    assert nat.rhs_info(rhs.diffusion_reaction.rhs_id)['kind'] == 1


def test_invalid_descriptors_are_rejected_on_the_host():
    from ozone_b200 import _native as nat
    pb = nat.OzProblem()
    pb.rhs_id = 99
    assert nat.lib.oz_tm_forward(C.byref(pb)) < 0 and b'rhs_id' in nat.lib.oz_last_error()
    pb.rhs_id, pb.num_stages, pb.N, pb.n, pb.num_inner = 0, 4, 8, 29, 5
    assert nat.lib.oz_tm_forward(C.byref(pb)) < 0 and b'divisible' in nat.lib.oz_last_error()   # time_marching_uniform.py:25-26
    pb.num_inner, pb.num_stages = 0, 9
    assert nat.lib.oz_tm_forward(C.byref(pb)) < 0
    assert nat.lib.oz_tm_adjoint_static(None) < 0


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, 'ozone_b200')
    for fn in os.listdir(pkg):
        if fn.endswith('.py'):
            assert 'oracle' not in open(os.path.join(pkg, fn)).read().replace('# oracle', ''), fn
