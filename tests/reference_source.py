"""Access to the reference's own source where it can be imported (test infrastructure).

``/root/reference`` exists in the build container only (never on the GPU box).  ``ozone/glms/**`` is plain
numpy and imports once a stand-in module named ``csdl_alpha`` is in ``sys.modules``; the right-hand side
callbacks of the examples are taken out of their files with ``ast`` (the files themselves build csdl
recorders at import time).  Nothing here is copied into the repository: the callers compare what they get
with this repo's data, or write derived VALUES (not source) into ``tests/golden``.
"""
import ast
import contextlib
import os
import sys
import types

REFERENCE = '/root/reference'
AVAILABLE = os.path.isdir(os.path.join(REFERENCE, 'ozone', 'glms'))

RHS_SOURCES = {   # problem -> (file under REFERENCE, function name)
    'van_der_pol': ('ozone/paper_examples/van_der_pol/main_model.py', 'ode_func'),
    'lunar_ascent': ('ozone/paper_examples/ascent_system/main_model.py', 'ode_func'),
    'lotka_volterra': ('examples/simple_examples/lotka_volterra.py', 'ode_function'),
    'pde_control': ('ozone/paper_examples/pde_control/pde_control.py', 'ode_function'),
}


def _stub_csdl():
    stub = types.ModuleType('csdl_alpha')

    def missing(name):
        if name.startswith('__'):
            raise AttributeError(name)
        return type(name, (), {})
    stub.__getattr__ = missing
    return stub


@contextlib.contextmanager
def reference_glms():
    """Yields the reference's ``ozone.glms.GLMs`` module; leaves ``sys.modules`` / ``sys.path`` as it found them."""
    before = dict(sys.modules)
    sys.modules['csdl_alpha'] = _stub_csdl()
    sys.path.insert(0, REFERENCE)
    try:
        import importlib
        yield importlib.import_module('ozone.glms.GLMs')
    finally:
        sys.path.remove(REFERENCE)
        for k in list(sys.modules):
            if k not in before:
                del sys.modules[k]


def reference_callback(problem, namespace=None):
    """The reference's right-hand side callback for `problem`, compiled from its own source text.  `namespace`: names the
    function body expects from its module (``csdl`` for the callbacks that call csdl functions)."""
    import numpy as np
    path, name = RHS_SOURCES[problem]
    tree = ast.parse(open(os.path.join(REFERENCE, path)).read())
    node = next(n for n in ast.walk(tree) if isinstance(n, ast.FunctionDef) and n.name == name)
    for a in node.args.args:
        a.annotation = None               # ozone.ODEVars / csdl.Variable hints: names that need not exist here
    mod = ast.Module(body=[node], type_ignores=[])
    ns = {'np': np}
    ns.update(namespace or {})
    exec(compile(ast.fix_missing_locations(mod), os.path.join(REFERENCE, path), 'exec'), ns)
    return ns[name]
