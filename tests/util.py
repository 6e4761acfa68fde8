import numpy as np


def relerr(a, b):
    """Norm-wise relative error of SURVEY.md section 8d: max|a-b| / max(max|b|, 1e-300)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300))


TOL_F64 = 1e-9  # north_star: FP64 state and covariance within 1e-9 relative

# TF32 variant of the large-map UKF's sigma-point products (tcgen05 kind::tf32, FP32 accumulation in TMEM), stated
# separately from the FP64 bar. (X, P) norm-wise relative to the FP64 path of the same library.
TOL_TF32_GEMM = 2e-3      # one pass: operands rounded to 11 significant bits
TOL_TF32X3_GEMM = 1e-4    # hi/lo split: what remains is the FP32 accumulation in TMEM (5e-7 at K = 32, 3e-5 at K = 4107)
TOL_TF32 = (1e-3, 1e-2)
TOL_TF32X3 = (5e-6, 2e-4)  # measured at n = 131 / 515 / 2051: X 7e-7 / 3e-7 / 6e-7, P 2e-5 / 6e-5 / 5e-5 (TF32: X 3e-5, P 1.3e-3)
