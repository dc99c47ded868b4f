#!/usr/bin/env python
"""Write tests/golden/reference_tableaux.json from the reference's own ``ozone/glms`` module.

Run in the build container (``/root/reference`` present):  python tools/make_reference_goldens.py
Values are stored as C99 hex floats, so the fixture is exact.  Only derived values are written, no source."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import reference_source as RS   # noqa: E402

out = {}
with RS.reference_glms() as G:
    for name in sorted(G.method_classes):
        m = G.get_integration_method(name)
        A = np.asarray(m.A, dtype=np.float64)
        B = np.asarray(m.B, dtype=np.float64).reshape(-1)
        out[name] = {'A': [[float(x).hex() for x in row] for row in A], 'B': [float(x).hex() for x in B],
                     'explicit': bool(m.explicit)}
path = os.path.join(ROOT, 'tests', 'golden', 'reference_tableaux.json')
json.dump(out, open(path, 'w'), indent=1, sort_keys=True)
print('wrote', path, len(out), 'tableaux')
