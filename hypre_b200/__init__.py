"""hypre_b200 -- B200 (sm_100a) solve phase for hypre's BoomerAMG-PCG.

The product is two shared libraries under ``hypre_b200/lib``:

* ``libb200amg.so``   -- hand-written CUDA kernels behind the C ABI of ``include/b200amg.h``
* ``libHYPRE_b200.so`` -- the hypre-facing binding (``host/hypre_glue.c``) that interposes
  ``hypre_PCGSolve`` / ``HYPRE_ParCSRMatrixMatvec`` on top of ``libHYPRE_host.so``, the reference's host
  library built from the sources where they lie (it keeps doing the AMG *setup*, as the north star
  prescribes).  Nothing under ``hypre_b200/lib`` links or loads anything under ``oracle/``.

This Python package is only plumbing for tests and ``bench.py``: ctypes bindings that drive
the HYPRE C API the way ``src/test/ij.c`` does.  There is no CPU fallback: if the CUDA
library cannot be loaded every import of :mod:`hypre_b200.capi` fails loudly.
"""
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_DIR = os.path.join(ROOT, "hypre_b200", "lib")
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")

LIB_B200 = os.path.join(LIB_DIR, "libb200amg.so")
LIB_HYPRE_B200 = os.path.join(LIB_DIR, "libHYPRE_b200.so")
LIB_HYPRE_B200_MP = os.path.join(LIB_DIR, "libHYPRE_b200_mp.so")   # the binding over the MPI build of the reference
LIB_HYPRE_HOST = os.path.join(LIB_DIR, "libHYPRE_host.so")         # the reference's host library (setup phase), product build
LIB_BRIDGE = os.path.join(LIB_DIR, "libhypre_bridge.so")            # struct accessors for tests / bench.py (not product)
LIB_BRIDGE_MP = os.path.join(LIB_DIR, "libhypre_bridge_mp.so")
MPIRUN_SHIM = os.path.join(LIB_DIR, "mpirun_shim")                  # launcher of the native multi-rank mode
LIB_HYPRE_REF = os.path.join(REF_DIR, "libHYPRE_ref.so")            # the oracle's separately built reference (checker)
LIB_REF_BRIDGE = os.path.join(REF_DIR, "libref_bridge.so")          # the same accessors, the checker's own build
LIB_ORACLE = os.path.join(REF_DIR, "liboracle.so")
IJ_B200 = os.path.join(LIB_DIR, "ij_b200")
IJ_REF = os.path.join(REF_DIR, "ij_ref")

__all__ = ["ROOT", "LIB_DIR", "LIB_B200", "LIB_HYPRE_B200", "LIB_HYPRE_B200_MP", "LIB_HYPRE_REF", "LIB_ORACLE",
           "LIB_HYPRE_HOST", "LIB_BRIDGE", "LIB_BRIDGE_MP", "MPIRUN_SHIM",
           "IJ_B200", "IJ_REF"]
