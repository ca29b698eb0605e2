"""CPU oracle for the deconvolution-GMM hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``density_deconvolution_b200/`` imports
this package; only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may.

Parity status: PINNED.  ``oracle/gmm_oracle.py`` is checked in
``tests/test_oracle_golden.py`` against vectors produced by importing the
reference itself (``tests/golden/make_golden.py`` -> ``tests/golden/*.npz``) and
against scipy closed forms.
"""
from . import gmm_oracle  # noqa: F401
