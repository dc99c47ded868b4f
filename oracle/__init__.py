"""CPU oracle for the solve-and-differentiate path of LSDOlab/ozone  --  TEST INFRASTRUCTURE ONLY.

This package is a numpy (fp64) restatement of the reference's integrator algorithms.  It is the
checker the CUDA path is compared against; it is never the thing shipped or measured.  Only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it.  Nothing under ``ozone_b200/`` imports it.

Why a restatement and not the reference itself: every core module of the reference imports
``csdl_alpha`` (``ozone/odeproblem.py:2``), an unpinned, un-vendored git dependency
(``setup.py:36``: ``csdl_alpha @ git+https://github.com/LSDOlab/CSDL_alpha.git``, no tag) that is
not installed and not installable here.  All arithmetic of the reference (loops, Newton,
Gauss-Seidel, reverse-mode AD) lives in that dependency, so the oracle restates the *published
algorithm structure* recorded by the reference's own integrator files and pins it on the
reference's own known answers.

PARITY IS PINNED by (see ``tests/test_oracle_goldens.py``):
  G1  tutorial GL4 trajectory, 11 values, 8 digits    examples/simple_examples/tutorial.ipynb:161-171
  G2  test_1 (y' = -y)  sum + 2 Jacobian norms        ozone/tests/test_correctness/test_1_simple/utils.py:41-52
  G3  test_2 (3 states, 5 dyn. params, outputs)       ozone/tests/test_correctness/test_2_general/utils.py:127-162
  G4  test_3 (500 independent trajectories)           ozone/tests/test_correctness/test_3_large/utils.py:55-74
G3's two outputs and G4's five values are reproduced to fp64 round-off (they were generated with RK4).

Restated files (cited again at each function):
  ozone/integrators/time_marching.py          explicit stages, implicit Newton residual, update, outputs
  ozone/integrators/time_marching_uniform.py  checkpoint layout
  ozone/integrators/picard_iteration.py       fixed-point sweep
  ozone/integrators/collocation.py            constraints c1, c2
  ozone/collections.py:256-293                stage-abscissa interpolation of dynamic parameters
  ozone/collections.py:125-139, 204-209       initial guesses and scalers
  ozone/glms/**                               tableaux

Array conventions (all numpy fp64, batch axis LAST so that shared inputs broadcast):
  y    (d, N)        packed state of N independent trajectories
  h    (n,) | (n, N) step sizes
  p    (nt, P, N) | (nt, P, 1)   dynamic parameters at the nt = n + 1 grid points
  th   (S, N) | (S, 1)           static ("closure captured") parameters
"""
