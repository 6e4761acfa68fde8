"""TEST INFRASTRUCTURE ONLY -- independent numpy restatement of the reference filter math.

Written from the reference sources (paths relative to /root/reference/awesome_slam/), NOT from
oracle/*.hpp, so that a mistake shared by the C++ oracle and the CUDA kernels (same author) has a
second, differently-structured implementation to disagree with (last row).
Dense algebra goes through numpy/LAPACK (different summation order from the oracle's scalar
loops); scalar transcendental calls go through `math` (glibc), float cast sites through np.float32.

    EKF   src/ekf/ekf.cpp:49-311      UKF   src/ukf/ukf.cpp:49-392, src/ukf/ukf.h:56-82
    f, h  include/awesome_slam/common.h:46-90     tools.h:44-66     structures.h:85-112
"""
import math

import numpy as np

f32 = np.float32
PI = f32(3.141592654)            # config.h:39
MIN_DIST_THRESH = f32(0.5)       # config.h:43
MIN_LANDMARK_OCC = 10            # config.h:44
UKF_STD_A = f32(0.2)             # config.h:54
UKF_STD_YAW = f32(0.2)           # config.h:55
UKF_KP_ROBOT_POSE = f32(0.001)   # config.h:56
UKF_KP_LANDMARK_POSE = f32(1.0)  # config.h:57
UKF_KR = f32(0.2)                # config.h:58
UKF_KQ = f32(0.001)              # config.h:59
EKF_KP_ROBOT_POSE = f32(0.001)   # config.h:62
EKF_KR = f32(0.2)                # config.h:64
EKF_KQ = f32(0.001)              # config.h:65


def normalize_angle(theta):
    """tools.h:44-50: float in, fmodf, float out."""
    theta = f32(theta)
    two_pi = f32(2) * PI
    ret = f32(math.fmod(float(theta), float(two_pi)))  # fmodf is exact, so fmod on the widened floats equals it
    if ret > PI:
        ret = f32(ret - two_pi)
    if ret < -PI:
        ret = f32(ret + two_pi)
    return ret


def euler_distance(p1, p2):
    """tools.h:53-59; the points already hold float-rounded coordinates (structures.h:52)."""
    dx = f32(p1[0] - p2[0])
    dy = f32(p1[1] - p2[1])
    return f32(np.sqrt(f32(f32(dx * dx) + f32(dy * dy))))


def to_point(rng, brg, Z):
    """structures.h:104-111: evaluated in double, stored to float."""
    a = f32(Z[0] + float(rng) * math.cos(Z[2] + float(brg)))
    b = f32(Z[1] + float(rng) * math.sin(Z[2] + float(brg)))
    return (float(a), float(b))


def state_transition(N, point, vx, az, dt):
    """common.h:46-75; vx, az, dt are floats."""
    vx, az, dt = f32(vx), f32(az), f32(dt)
    P = np.array(point[:N], np.float64)
    th = float(point[2])
    if abs(float(az)) > 0.001:
        r = float(f32(vx / az))
        azdt = float(f32(az * dt))
        P[0] += r * (-math.sin(th) + math.sin(th + azdt))
        P[1] += r * (math.cos(th) - math.cos(th + azdt))
    else:
        vxdt = float(f32(vx * dt))
        P[0] += vxdt * math.cos(th)
        P[1] += vxdt * math.sin(th)
    P[2] += float(f32(az * dt))
    if len(point) > N:
        d = float(dt)
        P[0] += 0.5 * d * d * point[N] * math.cos(th)
        P[1] += 0.5 * d * d * point[N] * math.sin(th)
        P[2] += 0.5 * d * d * float(az)
    return P


def measurement(N, point):
    """common.h:78-90."""
    P = np.array(point, np.float64)
    for i in range(0, N - 3, 2):
        dx, dy = point[3 + i] - point[0], point[4 + i] - point[1]
        P[3 + i] = math.sqrt(dx * dx + dy * dy)
        P[4 + i] = math.atan2(dy, dx) - point[2]
    return P


def eigen_llt_lower(A):
    """Eigen 3.3.x LLT<MatrixXd, Lower>::compute + matrixL(), restated from its published algorithm
    (Eigen/src/Cholesky/LLT.h, llt_inplace<Scalar, Lower>): size < 32 runs `unblocked`; otherwise the
    right-looking `blocked` loop with blockSize = clamp((size/8)/16*16, 8, 128). On a non-positive pivot the
    factorisation stops and the matrix is left half-factored; matrixL() is its lower triangle.
    Returns (L, fail_k) with fail_k = -1 on success."""
    m = np.array(A, np.float64)
    size = m.shape[0]

    def unblocked(off, sz):
        for k in range(sz):
            kk = off + k
            x = m[kk, kk]
            if k > 0:
                x -= float(np.dot(m[kk, off:kk], m[kk, off:kk]))
            if x <= 0:
                return k
            x = math.sqrt(x)
            m[kk, kk] = x
            rs = sz - k - 1
            if rs > 0:
                if k > 0:
                    m[kk + 1:off + sz, kk] -= m[kk + 1:off + sz, off:kk] @ m[kk, off:kk]
                m[kk + 1:off + sz, kk] /= x
        return -1

    if size < 32:
        ret = unblocked(0, size)
    else:
        bs_ = size // 8
        bs_ = (bs_ // 16) * 16
        bs_ = min(max(bs_, 8), 128)
        ret = -1
        for k in range(0, size, bs_):
            bs = min(bs_, size - k)
            rs = size - k - bs
            r = unblocked(k, bs)
            if r >= 0:
                ret = k + r
                break
            if rs > 0:
                A11 = np.tril(m[k:k + bs, k:k + bs])
                # A21 <- A21 * A11^-T
                m[k + bs:, k:k + bs] = np.linalg.solve(A11, m[k + bs:, k:k + bs].T).T
                A21 = m[k + bs:, k:k + bs]
                upd = A21 @ A21.T
                il = np.tril_indices(rs)
                sub = m[k + bs:, k + bs:]
                sub[il] -= upd[il]
    return np.tril(m), ret


class _Front:
    """ekf.cpp:137-203,217-253 / ukf.cpp:113-180,184-220: association + wait-list (identical in both)."""

    def __init__(self, cap):
        self.N = 3
        self.cap = cap
        self.obs = []
        self.wait = []  # [range, bearing, count]
        self.assoc = []
        self.init_x = self.init_z = True

    def set_observations(self, rng, brg):
        self.init_z = False
        self.obs = [[f32(r), f32(b)] for r, b in zip(rng, brg)]

    def _update_wait(self, data, Z):
        if not self.wait:
            self.wait.append([data[0], data[1], 1])
            return
        l2 = to_point(data[0], data[1], Z)
        best, corr = None, 0
        for i, w in enumerate(self.wait):
            d = euler_distance(l2, to_point(w[0], w[1], Z))
            if best is None or d < best:
                best, corr = d, i
        if best < MIN_DIST_THRESH:
            self.wait[corr][2] += 1
        else:
            self.wait.append([data[0], data[1], 1])

    def associate(self, X, Z):
        self.assoc = []
        N = self.N
        for data in self.obs:
            data[1] = normalize_angle(data[1])
            if N == 3:
                self._update_wait(data, Z)
                self.assoc.append(-1)
                continue
            l2 = to_point(data[0], data[1], Z)
            best, corr = None, 0
            for j in range(0, N - 3, 2):
                l1 = (float(f32(X[3 + j])), float(f32(X[4 + j])))
                d = euler_distance(l2, l1)
                if best is None or d < best:
                    best, corr = d, j
            if best < MIN_DIST_THRESH:
                Z[3 + corr] = float(data[0])
                Z[4 + corr] = float(data[1])
                self.assoc.append(corr)
            else:
                self._update_wait(data, Z)
                self.assoc.append(-1)
        fresh = []
        for w in self.wait:
            if w[2] == MIN_LANDMARK_OCC:
                fresh.append((w[0], w[1]))
                w[2] += 1
        return fresh


def _grown(M, N, fill_diag):
    """conservativeResizeLike(Identity(N,N) * fill_diag): old block kept, new entries from `other`."""
    out = np.eye(N) * float(fill_diag)
    k = M.shape[0]
    out[:k, :k] = M
    return out


class EkfNp(_Front):
    def __init__(self, cap=30):
        super().__init__(cap)
        self.X = np.zeros(3)
        self.Z = np.zeros(3)
        self.Q = np.eye(3) * float(EKF_KQ)
        self.R = np.eye(3) * float(EKF_KR)
        self.P = np.eye(3) * float(EKF_KP_ROBOT_POSE)
        self.A = np.eye(3)
        self.H = np.eye(3)

    def _grow(self, fresh):
        cacheN = self.N
        N = cacheN + 2 * len(fresh)
        if N >= self.cap:  # ekf.cpp:263-268
            return
        self.N = N
        self.X = np.concatenate([self.X, np.zeros(N - cacheN)])
        self.Z = np.concatenate([self.Z, np.zeros(N - cacheN)])
        self.Q = _grown(self.Q, N, 0.0)
        self.A = _grown(self.A, N, 1.0)
        self.H = _grown(self.H, N, 1.0)
        self.P = _grown(self.P, N, UKF_KP_LANDMARK_POSE)  # ekf.cpp:277
        self.R = _grown(self.R, N, UKF_KR)                # ekf.cpp:278
        for i, (r, b) in enumerate(fresh):
            k = cacheN + 2 * i
            self.Z[k], self.Z[k + 1] = float(r), float(b)
            self.X[k] = self.Z[0] + self.Z[k] * math.cos(self.Z[2] + self.Z[k + 1])
            self.X[k + 1] = self.Z[1] + self.Z[k] * math.sin(self.Z[2] + self.Z[k + 1])

    def step(self, ox, oy, yaw, vx, az, dt):
        if self.init_z:
            return False
        self.Z[0], self.Z[1], self.Z[2] = ox, oy, float(f32(yaw))
        fresh = self.associate(self.X, self.Z)
        if fresh:
            self._grow(fresh)
        if vx and az:  # ekf.cpp:206-212 (message doubles)
            dth = float(f32(az * float(f32(dt))))
            r = float(f32(vx / az))
            self.A[0, 0] = r * (-math.cos(self.Z[2]) + math.cos(self.Z[2] + dth))
            self.A[1, 0] = r * (-math.sin(self.Z[2]) + math.sin(self.Z[2] + dth))
        if self.init_x:
            self.init_x = False
            self.X = self.Z.copy()
        self._slam(f32(vx), f32(az), f32(dt))
        return True

    def _slam(self, vx, az, dt):  # ekf.cpp:293-311
        N = self.N
        self.X = state_transition(N, self.X, vx, az, dt)
        self.X[2] = float(normalize_angle(self.X[2]))
        self.P = self.A @ self.P @ self.A.T + self.Q
        X, H = self.X, self.H
        for i in range(0, N - 3, 2):  # ekf.cpp:117-134
            hyp = f32((X[3 + i] - X[0]) ** 2 + (X[4 + i] - X[1]) ** 2)
            dist = float(f32(np.sqrt(hyp)))
            hyp = float(hyp)
            H[3 + i, 0] = (-X[3 + i] + X[0]) / dist
            H[4 + i, 0] = -(-X[4 + i] + X[1]) / hyp
            H[3 + i, 1] = (-X[4 + i] + X[1]) / dist
            H[4 + i, 1] = (-X[3 + i] + X[0]) / hyp
            H[4 + i, 2] = -1
            H[3 + i, 3 + i] = -(-X[3 + i] + X[0]) / dist
            H[3 + i, 4 + i] = -(-X[4 + i] + X[1]) / dist
            H[4 + i, 3 + i] = (-X[4 + i] + X[1]) / hyp
            H[4 + i, 4 + i] = -(-X[3 + i] + X[0]) / hyp
        S = H @ self.P @ H.T + self.R
        K = self.P @ H.T @ np.linalg.inv(S)
        Y = self.Z - measurement(N, X)
        for j in range(0, N - 1, 2):
            Y[2 + j] = float(normalize_angle(Y[2 + j]))
        self.X = X + K @ Y
        self.P = (np.eye(N) - K @ H) @ self.P


class UkfNp(_Front):
    def __init__(self, cap=30):
        super().__init__(cap)
        self.X = np.zeros(3)
        self.Z = np.zeros(3)
        self.Q = np.eye(3) * float(UKF_KQ)
        self.R = np.eye(3) * float(UKF_KR)
        self.P = np.eye(3) * float(UKF_KP_ROBOT_POSE)
        self.llt_fail = -1
        self._weights(3)

    def _weights(self, N):  # ukf.h:73-81
        self.lam = f32(3.0 - (N + 2))
        denom = f32(f32(self.lam + f32(N)) + f32(2))
        weight = f32(0.5 / float(denom))
        self.w = np.full(2 * N + 5, float(weight))
        self.w[0] = float(f32(self.lam / denom))

    def _grow(self, fresh):  # ukf.cpp:222-257
        cacheN = self.N
        N = cacheN + 2 * len(fresh)
        if N >= self.cap:
            return
        self.N = N
        self.X = np.concatenate([self.X, np.zeros(N - cacheN)])
        self.Z = np.concatenate([self.Z, np.zeros(N - cacheN)])
        self.Q = _grown(self.Q, N, 0.0)
        self.P = _grown(self.P, N, UKF_KP_LANDMARK_POSE)
        self.R = _grown(self.R, N, UKF_KR)
        self._weights(N)
        for i, (r, b) in enumerate(fresh):
            k = cacheN + 2 * i
            self.Z[k], self.Z[k + 1] = float(r), float(b)
            self.X[k] = self.Z[0] + self.Z[k] * math.cos(self.Z[2] + self.Z[k + 1])
            self.X[k + 1] = self.Z[1] + self.Z[k] * math.sin(self.Z[2] + self.Z[k + 1])

    def step(self, ox, oy, yaw, vx, az, dt):
        if self.init_z:
            return False
        self.Z[0], self.Z[1], self.Z[2] = ox, oy, float(f32(yaw))
        fresh = self.associate(self.X, self.Z)
        if fresh:
            self._grow(fresh)
        if self.init_x:
            self.init_x = False
            self.X = self.Z.copy()
        self.slam(f32(vx), f32(az), f32(dt))
        return True

    def _wrap(self, v):
        for j in range(0, self.N - 1, 2):
            v[2 + j] = float(normalize_angle(v[2 + j]))
        return v

    def slam(self, vx, az, dt):  # ukf.cpp:260-392
        N = self.N
        NA, NS = N + 2, 2 * N + 5
        Xaug = np.concatenate([self.X, [0.0, 0.0]])
        Paug = np.zeros((NA, NA))
        Paug[:N, :N] = self.P
        Paug[N, N] = float(f32(UKF_STD_A * UKF_STD_A))
        Paug[N + 1, N + 1] = float(f32(UKF_STD_YAW * UKF_STD_YAW))
        L, self.llt_fail = eigen_llt_lower(Paug)
        w = float(f32(np.sqrt(f32(f32(self.lam + f32(N)) + f32(2)))))
        Xs = np.empty((NA, NS))
        Xs[:, 0] = Xaug
        Xs[:, 1:NA + 1] = Xaug[:, None] + w * L
        Xs[:, NA + 1:] = Xaug[:, None] - w * L
        Xp = np.empty((N, NS))
        for i in range(NS):
            Xp[:, i] = state_transition(N, Xs[:, i], vx, az, dt)
            Xp[2, i] = float(normalize_angle(Xp[2, i]))
        X = np.zeros(N)
        for i in range(NS):  # sequential accumulation, as written (ukf.cpp:300-304)
            X = X + self.w[i] * Xp[:, i]
        D = Xp - X[:, None]
        for i in range(NS):
            D[2, i] = float(normalize_angle(D[2, i]))
        P = np.zeros((N, N))
        for i in range(NS):
            P += np.outer(self.w[i] * D[:, i], D[:, i])
        P = P + self.Q
        Zs = np.empty((N, NS))
        for i in range(NS):
            Zs[:, i] = measurement(N, Xp[:, i])
        Zp = np.zeros(N)
        for i in range(NS):
            Zp = Zp + self.w[i] * Zs[:, i]
        self._wrap(Zp)
        DZ = Zs - Zp[:, None]
        for i in range(NS):
            self._wrap(DZ[:, i])
        S = np.zeros((N, N))
        Tc = np.zeros((N, N))
        for i in range(NS):
            S += np.outer(self.w[i] * DZ[:, i], DZ[:, i])
            Tc += np.outer(self.w[i] * D[:, i], DZ[:, i])
        S = S + self.R
        K = Tc @ np.linalg.inv(S)
        dz = self._wrap(self.Z - Zp)
        self.X = X + K @ dz
        self.P = P - K @ S @ K.T


# ---------------------------------------------------------------------------------------------------------------
# Vectorised form of one UKF message for LARGE maps (cfg5: n = 2051 finishes in seconds instead of minutes). Same
# formulas as UkfNp (ukf.cpp:113-180, 260-392) with the per-sigma-point loops written as matrix operations; the
# map does not grow (unmatched observations are reported as -1 and dropped). tests/test_oracle.py checks it against
# the loop form above at a size both can run.
def normalize_angle_vec(theta):
    t = np.asarray(theta, np.float64).astype(np.float32)
    two_pi = f32(2) * PI
    r = np.fmod(t, two_pi).astype(np.float32)
    r = np.where(r > PI, (r - two_pi).astype(np.float32), r)
    r = np.where(r < -PI, (r + two_pi).astype(np.float32), r)
    return r.astype(np.float64)


def ukf_large_step(X, Z, P, ox, oy, yaw, vx, az, dt, rng, brg):
    """Returns (X', Z', P', assoc, llt_fail). P indexed [i, j]."""
    X = np.array(X, np.float64)
    Z = np.array(Z, np.float64)
    N = X.size
    Z[0], Z[1], Z[2] = ox, oy, float(f32(yaw))
    lmx32 = X[3::2].astype(np.float32).astype(np.float64)
    lmy32 = X[4::2].astype(np.float32).astype(np.float64)
    assoc = []
    for r, b in zip(rng, brg):
        r, b = f32(r), normalize_angle(b)
        px, py = to_point(r, b, Z)
        dx = (px - lmx32).astype(np.float32)
        dy = (py - lmy32).astype(np.float32)
        d = np.sqrt((dx * dx).astype(np.float32) + (dy * dy).astype(np.float32)).astype(np.float32)
        c = int(np.argmin(d))  # first minimum, as the strict `<` of ukf.cpp:144
        if d[c] < MIN_DIST_THRESH:
            Z[3 + 2 * c], Z[4 + 2 * c] = float(r), float(b)
            assoc.append(2 * c)
        else:
            assoc.append(-1)
    vx, az, dt = f32(vx), f32(az), f32(dt)
    lam = f32(3.0 - (N + 2))
    denom = f32(f32(lam + f32(N)) + f32(2))
    wts = np.full(2 * N + 5, float(f32(0.5 / float(denom))))
    wts[0] = float(f32(lam / denom))
    NA, NS = N + 2, 2 * N + 5
    Paug = np.zeros((NA, NA))
    Paug[:N, :N] = P
    Paug[N, N] = float(f32(UKF_STD_A * UKF_STD_A))
    Paug[N + 1, N + 1] = float(f32(UKF_STD_YAW * UKF_STD_YAW))
    try:
        L = np.linalg.cholesky(Paug)
        fail = -1
    except np.linalg.LinAlgError:
        L, fail = eigen_llt_lower(Paug)
    w = float(f32(np.sqrt(denom)))
    Xaug = np.concatenate([X, [0.0, 0.0]])
    Xs = np.empty((NA, NS))
    Xs[:, 0] = Xaug
    Xs[:, 1:NA + 1] = Xaug[:, None] + w * L
    Xs[:, NA + 1:] = Xaug[:, None] - w * L
    Xp = Xs[:N].copy()
    for i in range(NS):
        col = np.concatenate([Xs[:3, i], np.zeros(N - 3), Xs[N:, i]])  # f only reads rows 0..2 and N
        Xp[:3, i] = state_transition(N, col, vx, az, dt)[:3]
    Xp[2] = normalize_angle_vec(Xp[2])
    mean = Xp @ wts
    D = Xp - mean[:, None]
    D[2] = normalize_angle_vec(D[2])
    Q = np.zeros(N)
    Q[:3] = float(UKF_KQ)
    Pp = (D * wts) @ D.T + np.diag(Q)
    Zs = Xp.copy()
    dx = Xp[3::2] - Xp[0]
    dy = Xp[4::2] - Xp[1]
    Zs[3::2] = np.sqrt(dx * dx + dy * dy)
    Zs[4::2] = np.arctan2(dy, dx) - Xp[2]
    Zp = Zs @ wts
    Zp[2::2] = normalize_angle_vec(Zp[2::2])
    DZ = Zs - Zp[:, None]
    DZ[2::2] = normalize_angle_vec(DZ[2::2])
    S = (DZ * wts) @ DZ.T + np.eye(N) * float(UKF_KR)
    Tc = (D * wts) @ DZ.T
    K = np.linalg.solve(S.T, Tc.T).T
    dz = Z - Zp
    dz[2::2] = normalize_angle_vec(dz[2::2])
    Xn = mean + K @ dz
    Pn = Pp - K @ S @ K.T
    return Xn, Z, Pn, np.array(assoc, np.int32), fail


# ------------------------------------------------------------------------------------------------------------------
# Laser front end: independent numpy restatement of sensor_landmark.cpp, used only to cross-check
# oracle/laser_front.hpp. numpy.linalg.svd / eigh / solve stand in for Eigen's JacobiSVD / SelfAdjointEigenSolver /
# colPivHouseholderQr; float32 arithmetic is done with numpy float32 scalars.
# ------------------------------------------------------------------------------------------------------------------
_F = np.float32


def _laser_tables():
    th = np.arange(360, dtype=np.float32)
    ang = (_F(0.01745329251) * th).astype(np.float32)
    return np.sin(ang).astype(np.float32), np.cos(ang).astype(np.float32)


def _ldist(p, q):
    dx = _F(np.float64(p[0]) - np.float64(q[0]))
    dy = _F(np.float64(p[1]) - np.float64(q[1]))
    return np.sqrt(_F(_F(dx * dx) + _F(dy * dy)))


def laser_classify(pts):
    """sensor_landmark.cpp:147-186"""
    p1, p2 = pts[0], pts[-1]
    ang = []
    s = _F(0)
    with np.errstate(all="ignore"):
        for p in pts[1:-1]:
            c, a, b = _ldist(p1, p2), _ldist(p1, p), _ldist(p2, p)
            num = _F(_F(_F(a * a) + _F(b * b)) - _F(c * c))
            v = np.arccos(_F(num / _F(_F(2) * a * b)))
            ang.append(_F(v))
            s = _F(s + _F(v))
        mean = _F(s / _F(len(pts) - 2))
        var = _F(0)
        for v in ang:
            var = _F(np.float64(var) + np.float64(_F(v - mean)) ** 2)
        sd = _F(np.sqrt(np.float64(var) / 5.0))
    return bool(sd < _F(0.4) and mean > _F(1.5) and mean < _F(3.0))


def laser_fit(pts):
    """sensor_landmark.cpp:192-298 + structures.h:126-133 -> (range, bearing) float32"""
    n = len(pts)
    xm = _F(0)
    ym = _F(0)
    for p in pts:
        xm = _F(xm + _F(p[0]))
        ym = _F(ym + _F(p[1]))
    xm = _F(xm / _F(n))
    ym = _F(ym / _F(n))
    q = [(_F(_F(p[0]) - xm), _F(_F(p[1]) - ym)) for p in pts]
    z = [_F(_F(a * a) + _F(b * b)) for a, b in q]
    zm = _F(0)
    for v in z:
        zm = _F(zm + v)
    zm = _F(zm / _F(n))
    Z = np.array([[np.float64(z[i]), np.float64(q[i][0]), np.float64(q[i][1]), 1.0] for i in range(n)])
    Hinv = np.array([[0, 0, 0, 0.5], [0, 1, 0, 0], [0, 0, 1, 0], [0.5, 0, 0, np.float64(_F(_F(-2) * zm))]], dtype=np.float64)
    U, sv, Vt = np.linalg.svd(Z, full_matrices=False)
    V = Vt.T
    if sv[3] > 10e-12:
        Y = V @ np.diag(sv) @ V.T
        Q = Y @ Hinv @ Y
        ev, E = np.linalg.eigh(Q)
        sid, sev = 0, _F(99999)
        for i in range(4):
            if ev[i] > 0 and ev[i] < np.float64(sev):
                sid, sev = i, _F(ev[i])
        A = np.linalg.solve(Y, E[:, sid])
    else:
        A = V[:, 3]
    a = _F(-A[1] / (2 * A[0]))
    b = _F(-A[2] / (2 * A[0]))
    cx, cy = np.float64(_F(a + xm)), np.float64(_F(b + ym))
    return _F(np.sqrt(cx * cx + cy * cy)), _F(np.arctan2(cy, cx))


def laser_extract(ranges):
    """sensor_landmark.cpp:59-133 for one scan of 360 float32 ranges -> list of (range, bearing), or None where
    the reference's assert(theta < 359) fires."""
    sin_map, cos_map = _laser_tables()
    with np.errstate(all="ignore"):
        r32 = ranges.astype(np.float32)
        # wrap test (:63-64): rays 0 and 359 with their own table entries
        w0 = (_F(r32[0] * cos_map[0]), _F(r32[0] * sin_map[0]))
        w359 = (_F(r32[359] * cos_map[359]), _F(r32[359] * sin_map[359]))
        if _ldist(w0, w359) < _F(0.5):
            return None
        # (:87,98) bearing2pose(theta, ranges[theta++]) as GCC evaluates it (right to left): ray t uses table entry
        # t + 1; entry 360 lies one past the table and reads 0.0f (see oracle/laser_front.hpp, pinned by oracle/_ref)
        cs = np.concatenate([cos_map[1:], np.zeros(1, np.float32)]).astype(np.float32)
        sn = np.concatenate([sin_map[1:], np.zeros(1, np.float32)]).astype(np.float32)
        px = (r32 * cs).astype(np.float32)
        py = (r32 * sn).astype(np.float32)
        pts = list(zip(px, py))
        out, cluster, p1 = [], [pts[0]], pts[0]
        for th in range(1, 360):
            p2 = pts[th]
            if _ldist(p1, p2) < _F(0.5):
                cluster.append(p2)
            elif len(cluster) > 3 and laser_classify(cluster):
                out.append(laser_fit(cluster))
                cluster = []
            else:
                cluster = []
            p1 = p2
    return out
