"""trilinos_b200 — B200-native Block GMRES hot path behind Belos' plugin surface.

Layout: csrc/ (sm_100a kernels + C-ABI), lib/ (built .so, git-ignored), core.py (ctypes mirror
of Belos::MultiVec / Belos::Operator), solver.py (unmodified Belos::BlockGmresSolMgr driver).
"""
from ._lib import B200Error, EXPORTED, LIB_PATH, SOLVER_LIB_PATH  # noqa: F401
from .core import (read_hb, read_matrix_market, Chebyshev, ChebyParams, Context, CrsOperator, MultiVec, NORM_INF, NORM_ONE, NORM_TWO, build_colmap,  # noqa: F401
                   galeri, linear_partition)
from .solver import (BlockCGSolMgr, BlockGmresSolMgr, PseudoBlockCGSolMgr, PseudoBlockGmresSolMgr,
                     SOLVER_MANAGERS, mvop_test, ortho_test)  # noqa: F401,E402
from . import dist  # noqa: F401,E402
