"""CPU oracle for the ISDF hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import it, and only as the checker or the timed CPU
baseline.  The product (``pyscf_isdf_b200``) never imports this package and
fails loudly when its CUDA library is missing.

Pinning status (see DESIGN.md "Oracle"):

* ``eval_ao`` (stage 1), ``fft/ifft/get_coulG`` (stage 4) and the exact-grid
  ``fftdf_get_j/get_k`` limit (stage 5) restate reference code and are pinned
  by the reference's own known-answer tests (``tests/test_oracle_golden.py``:
  coulG fingerprints pyscf/pbc/tools/test/test_pbc.py:54-80, FFTDF J/K
  fingerprints pyscf/pbc/df/test/test_fft.py:642-645, d-shell AO fingerprint
  pyscf/pbc/gto/test/test_cell.py:500-504).
* ``select_ip / fit_aux / build_W / isdf_get_jk`` (stages 2, 3, ISDF form of
  4 and 5): the mounted reference contains NO ISDF code, so these follow the
  textbook formulation with the reference's FFTDF normalisation and the pivot
  rule of pyscf/lib/scipy_helper.py:71-110 -- **parity unpinned** for these.
"""
from .pbc_tools import fp, fft, ifft, get_coulG          # noqa: F401
from .eval_ao import eval_ao                               # noqa: F401
from .fftdf_jk import fftdf_get_j, fftdf_get_k, fftdf_get_jk   # noqa: F401
from .isdf_ref import (select_ip, fit_aux, build_V, build_W, isdf_get_jk,    # noqa: F401
                       isdf_build)
