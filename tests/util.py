import numpy as np


def relerr(a, b):
    """Norm-wise relative error of SURVEY.md section 8d: max|a-b| / max(max|b|, 1e-300)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300))


TOL_F64 = 1e-9  # north_star: FP64 state and covariance within 1e-9 relative
