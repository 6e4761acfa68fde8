// TEST INFRASTRUCTURE ONLY -- C entry points of the CPU oracle (see oracle/README.md).
// Loaded with ctypes from tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs.
// The product library (awesomeslam_b200/csrc) never links or calls this.
#include <chrono>
#include <cstring>
#include <memory>

#include "laser_front.hpp"
#include "slam_filters.hpp"

#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace {

struct Handle {
    int kind;  // 0 EKF, 1 UKF
    std::unique_ptr<EkfOracle> e;
    std::unique_ptr<UkfOracle> u;
    Frontend& fe() { return kind == 0 ? e->fe : u->fe; }
    Mat& X() { return kind == 0 ? e->X : u->X; }
    Mat& Z() { return kind == 0 ? e->Z : u->Z; }
    Mat& P() { return kind == 0 ? e->P : u->P; }
};

// Field order mirrors include/aslam_c.h's aslam_config on purpose (declared separately so
// the oracle has no dependency on the product headers).
struct orc_config {
    float pi, min_dist_thresh;
    int32_t min_landmark_occ, max_state_dim;
    float ukf_std_a, ukf_std_yaw, ukf_kp_robot_pose, ukf_kp_landmark_pose, ukf_kr, ukf_kq;
    float ekf_kp_robot_pose, ekf_kp_landmark_pose, ekf_kr, ekf_kq;
};

Config to_cfg(const orc_config* c) {
    Config k;
    if (!c) return k;
    k.pi = c->pi;
    k.min_dist_thresh = c->min_dist_thresh;
    k.min_landmark_occ = c->min_landmark_occ;
    k.max_state_dim = c->max_state_dim;
    k.ukf_std_a = c->ukf_std_a;
    k.ukf_std_yaw = c->ukf_std_yaw;
    k.ukf_kp_robot_pose = c->ukf_kp_robot_pose;
    k.ukf_kp_landmark_pose = c->ukf_kp_landmark_pose;
    k.ukf_kr = c->ukf_kr;
    k.ukf_kq = c->ukf_kq;
    k.ekf_kp_robot_pose = c->ekf_kp_robot_pose;
    k.ekf_kp_landmark_pose = c->ekf_kp_landmark_pose;
    k.ekf_kr = c->ekf_kr;
    k.ekf_kq = c->ekf_kq;
    return k;
}

}  // namespace

extern "C" {

void orc_config_default(orc_config* c) {
    Config k;
    c->pi = k.pi;
    c->min_dist_thresh = k.min_dist_thresh;
    c->min_landmark_occ = k.min_landmark_occ;
    c->max_state_dim = k.max_state_dim;
    c->ukf_std_a = k.ukf_std_a;
    c->ukf_std_yaw = k.ukf_std_yaw;
    c->ukf_kp_robot_pose = k.ukf_kp_robot_pose;
    c->ukf_kp_landmark_pose = k.ukf_kp_landmark_pose;
    c->ukf_kr = k.ukf_kr;
    c->ukf_kq = k.ukf_kq;
    c->ekf_kp_robot_pose = k.ekf_kp_robot_pose;
    c->ekf_kp_landmark_pose = k.ekf_kp_landmark_pose;
    c->ekf_kr = k.ekf_kr;
    c->ekf_kq = k.ekf_kq;
}

void* orc_create(int kind, const orc_config* c) {
    Handle* h = new Handle;
    h->kind = kind;
    if (kind == 0)
        h->e.reset(new EkfOracle(to_cfg(c)));
    else
        h->u.reset(new UkfOracle(to_cfg(c)));
    return h;
}

void orc_destroy(void* p) { delete (Handle*)p; }

void orc_set_observations(void* p, const float* range, const float* bearing, uint32_t n) {
    Handle* h = (Handle*)p;
    if (h->kind == 0)
        h->e->set_observations(range, bearing, n);
    else
        h->u->set_observations(range, bearing, n);
}

int orc_step(void* p, double x, double y, float yaw, double vx, double az, float dt) {
    Handle* h = (Handle*)p;
    return h->kind == 0 ? (int)h->e->step(x, y, yaw, vx, az, dt) : (int)h->u->step(x, y, yaw, vx, az, dt);
}

uint32_t orc_dim(void* p) { return ((Handle*)p)->fe().N; }

// X, Z: N doubles; P: N*N doubles column-major (Eigen's layout). Any pointer may be null.
void orc_get(void* p, double* X, double* Z, double* P) {
    Handle* h = (Handle*)p;
    const size_t N = h->fe().N;
    if (X) std::memcpy(X, h->X().d.data(), N * sizeof(double));
    if (Z) std::memcpy(Z, h->Z().d.data(), N * sizeof(double));
    if (P) std::memcpy(P, h->P().d.data(), N * N * sizeof(double));
}

// EKF only: the two information-bearing entries of A (ekf.cpp:210-211).
void orc_get_a(void* p, double* a00, double* a10) {
    Handle* h = (Handle*)p;
    *a00 = h->e->A(0, 0);
    *a10 = h->e->A(1, 0);
}

uint32_t orc_get_assoc(void* p, int32_t* out, uint32_t cap) {
    Handle* h = (Handle*)p;
    const auto& a = h->fe().assoc;
    for (uint32_t i = 0; i < a.size() && i < cap; ++i) out[i] = a[i];
    return (uint32_t)a.size();
}

uint32_t orc_get_wait(void* p, float* range, float* bearing, uint32_t* count, uint32_t cap) {
    Handle* h = (Handle*)p;
    const auto& w = h->fe().wait;
    for (uint32_t i = 0; i < w.size() && i < cap; ++i) {
        range[i] = w[i].first.range;
        bearing[i] = w[i].first.bearing;
        count[i] = w[i].second;
    }
    return (uint32_t)w.size();
}

int orc_llt_fail(void* p) {
    Handle* h = (Handle*)p;
    return h->kind == 1 ? h->u->log.llt_fail_k : -1;
}

int orc_cap_rejected(void* p) { return ((Handle*)p)->fe().cap_rejected ? 1 : 0; }

// Start from a supplied state (cfg 4/5): sizes every matrix exactly as N/2-1 promotions
// would have (ekf.cpp:271-278 / ukf.cpp:238-245), then overwrites X, Z, P.
void orc_load_state(void* p, uint32_t N, const double* X, const double* Z, const double* P) {
    Handle* h = (Handle*)p;
    const int n = (int)N;
    Frontend& fe = h->fe();
    const Config& c = fe.cfg;
    fe.N = N;
    Mat x(n, 1), z(n, 1), pm(n, n);
    std::memcpy(x.d.data(), X, sizeof(double) * n);
    std::memcpy(z.d.data(), Z, sizeof(double) * n);
    std::memcpy(pm.d.data(), P, sizeof(double) * n * n);
    Mat Q = Mat::zero(n, n);
    if (h->kind == 0) {
        EkfOracle& e = *h->e;
        e.init_x = false;
        Q(0, 0) = Q(1, 1) = Q(2, 2) = c.ekf_kq;
        e.X = x; e.Z = z; e.P = pm; e.Q = Q;
        e.I = Mat::identity(n, n);
        e.A = Mat::identity(n, n);
        e.H = Mat::identity(n, n);
        e.R = scale(Mat::identity(n, n), (double)c.ukf_kr);
        for (int i = 0; i < 3; ++i) e.R(i, i) = c.ekf_kr;
    } else {
        UkfOracle& u = *h->u;
        u.init_x = false;
        Q(0, 0) = Q(1, 1) = Q(2, 2) = c.ukf_kq;
        u.X = x; u.Z = z; u.P = pm; u.Q = Q;
        u.R = scale(Mat::identity(n, n), (double)c.ukf_kr);
        u.update_weights(N);
    }
}

// Runs B independent filters for T steps from a step-major replay ([T][B] odometry,
// [T][B][L] observations, every filter sees L observations per step) on `nthreads` host
// threads (OpenMP over filters). Outputs are padded to `ld`: X [B][ld], P [B][ld*ld]
// column-major with leading dimension ld, N [B]. Returns wall seconds of the filter loop.
double orc_run_batch(int kind, const orc_config* c, uint32_t B, uint32_t T, uint32_t L,
                     const double* ox, const double* oy, const float* yaw, const double* vx,
                     const double* az, const float* dt, const float* range, const float* bearing,
                     int nthreads, uint32_t ld, double* outX, double* outP, uint32_t* outN) {
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t b = 0; b < (int64_t)B; ++b) {
        Handle h;
        h.kind = kind;
        if (kind == 0)
            h.e.reset(new EkfOracle(to_cfg(c)));
        else
            h.u.reset(new UkfOracle(to_cfg(c)));
        for (uint32_t t = 0; t < T; ++t) {
            const size_t i = (size_t)t * B + b;
            orc_set_observations(&h, range + i * L, bearing + i * L, L);
            orc_step(&h, ox[i], oy[i], yaw[i], vx[i], az[i], dt[i]);
        }
        const uint32_t N = h.fe().N;
        if (outN) outN[b] = N;
        if (outX)
            for (uint32_t k = 0; k < N && k < ld; ++k) outX[(size_t)b * ld + k] = h.X()(k);
        if (outP)
            for (uint32_t j = 0; j < N && j < ld; ++j)
                for (uint32_t k = 0; k < N && k < ld; ++k)
                    outP[(size_t)b * ld * ld + (size_t)j * ld + k] = h.P()(k, j);
    }
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// ---------------------------------------------------------------------------------------
// Large-map EKF: row-subset joint update.
// Same formulas as ekf.cpp:293-311 with H restricted to the rows of the pose (0,1,2) and of
// the landmarks observed this step; with every landmark observed it IS ekf.cpp:300-310.
//   X [n], P [n*n] column-major, both updated in place.
//   a00/a10: the persistent A(0,0), A(1,0) (ekf.cpp:210-211), updated in place.
//   obs: m (range, bearing) floats; assoc_out[m] receives the landmark slot (j, even, as
//   in ekf.cpp:170) or -1 when no landmark is within MIN_DIST_THRESH (those observations
//   are not used; growth is outside the large-map contract).
// Returns the number of H rows used (3 + 2 * distinct observed landmarks).
// ---------------------------------------------------------------------------------------
int orc_ekf_large_step(const orc_config* cc, uint32_t n, double* X, double* P, double* a00, double* a10,
                       double ox, double oy, float yaw, double vx, double az, float dt,
                       const float* range, const float* bearing, uint32_t m, int32_t* assoc_out,
                       int nthreads) {
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    const Config c = to_cfg(cc);
    const size_t N = n;
    auto Pm = [&](size_t i, size_t j) -> double& { return P[j * N + i]; };

    // --- updateZandA front half (ekf.cpp:139-187) against the supplied map
    Mat Zp(3, 1);
    Zp(0) = ox; Zp(1) = oy; Zp(2) = yaw;
    std::vector<int> lm_obs(((n - 3) / 2), -1);  // landmark -> index of the last obs written to it
    std::vector<float> nb(m);
    for (uint32_t i = 0; i < m; ++i) {
        Laser d{range[i], normalize_angle(c, bearing[i])};
        nb[i] = d.bearing;
        int corr = 0;
        Pt l1((float)X[3], (float)X[4]);
        Pt l2 = to_point(d, Zp);
        float mindist = euler_distance(l2, l1);
        for (uint32_t j = 2; j < n - 3; j += 2) {
            l1 = Pt((float)X[3 + j], (float)X[4 + j]);
            float dist = euler_distance(l2, l1);
            if (dist < mindist) { corr = (int)j; mindist = dist; }
        }
        if (mindist < c.min_dist_thresh) {
            assoc_out[i] = corr;
            lm_obs[corr / 2] = (int)i;
        } else {
            assoc_out[i] = -1;
        }
    }
    if (vx && az) {  // ekf.cpp:206-212
        float delta_theta = (float)(az * (double)dt);
        float r = (float)(vx / az);
        *a00 = (double)r * (-std::cos(Zp(2)) + std::cos(Zp(2) + (double)delta_theta));
        *a10 = (double)r * (-std::sin(Zp(2)) + std::sin(Zp(2) + (double)delta_theta));
    }

    // --- predict (ekf.cpp:295-297). A = I except A(0,0), A(1,0).
    {
        Mat pose(3, 1);
        pose(0) = X[0]; pose(1) = X[1]; pose(2) = X[2];
        Mat np = state_transition(3, pose, (float)vx, (float)az, dt);
        X[0] = np(0); X[1] = np(1);
        X[2] = normalize_angle(c, (float)np(2));
    }
    const double A00 = *a00, A10 = *a10;
    // (A P): row0' = A00*row0 ; row1' = A10*row0 + row1     (k-ascending sum order of matmul)
    for (size_t j = 0; j < N; ++j) {
        const double p0 = Pm(0, j), p1 = Pm(1, j);
        Pm(0, j) = A00 * p0;
        Pm(1, j) = A10 * p0 + p1;
    }
    // (.) A^T: col0' = col0*A00 ; col1' = col0*A10 + col1
    for (size_t i = 0; i < N; ++i) {
        const double p0 = Pm(i, 0), p1 = Pm(i, 1);
        Pm(i, 0) = p0 * A00;
        Pm(i, 1) = p0 * A10 + p1;
    }
    Pm(0, 0) += (double)c.ekf_kq;
    Pm(1, 1) += (double)c.ekf_kq;
    Pm(2, 2) += (double)c.ekf_kq;

    // --- rows of H in use
    std::vector<uint32_t> rows = {0, 1, 2};
    std::vector<double> yv;  // innovation before wrap
    std::vector<double> rdiag = {(double)c.ekf_kr, (double)c.ekf_kr, (double)c.ekf_kr};
    yv.push_back(ox - X[0]);
    yv.push_back(oy - X[1]);
    yv.push_back((double)yaw - X[2]);
    struct HRow { uint32_t col[5]; double v[5]; int nnz; };
    std::vector<HRow> H;
    for (int k = 0; k < 3; ++k) { HRow h{}; h.nnz = 1; h.col[0] = (uint32_t)k; h.v[0] = 1.0; H.push_back(h); }
    for (uint32_t l = 0; l < (uint32_t)lm_obs.size(); ++l) {
        if (lm_obs[l] < 0) continue;
        const uint32_t i = 2 * l;
        // ekf.cpp:121-132
        float hyp = (float)(std::pow(X[3 + i] - X[0], 2) + std::pow(X[4 + i] - X[1], 2));
        float dist = std::sqrt(hyp);
        HRow hr{}, hb{};
        hr.nnz = 4;
        hr.col[0] = 0;     hr.v[0] = (-X[3 + i] + X[0]) / dist;
        hr.col[1] = 1;     hr.v[1] = (-X[4 + i] + X[1]) / dist;
        hr.col[2] = 3 + i; hr.v[2] = -(-X[3 + i] + X[0]) / dist;
        hr.col[3] = 4 + i; hr.v[3] = -(-X[4 + i] + X[1]) / dist;
        hb.nnz = 5;
        hb.col[0] = 0;     hb.v[0] = -(-X[4 + i] + X[1]) / hyp;
        hb.col[1] = 1;     hb.v[1] = (-X[3 + i] + X[0]) / hyp;
        hb.col[2] = 2;     hb.v[2] = -1;
        hb.col[3] = 3 + i; hb.v[3] = (-X[4 + i] + X[1]) / hyp;
        hb.col[4] = 4 + i; hb.v[4] = -(-X[3 + i] + X[0]) / hyp;
        H.push_back(hr);
        H.push_back(hb);
        rows.push_back(3 + i);
        rows.push_back(4 + i);
        rdiag.push_back((double)c.ukf_kr);
        rdiag.push_back((double)c.ukf_kr);
        // common.h:86-87 and ekf.cpp:302
        const double hr_pred = std::sqrt(std::pow(X[3 + i] - X[0], 2) + std::pow(X[4 + i] - X[1], 2));
        const double hb_pred = std::atan2(X[4 + i] - X[1], X[3 + i] - X[0]) - X[2];
        yv.push_back((double)range[lm_obs[l]] - hr_pred);
        yv.push_back((double)nb[lm_obs[l]] - hb_pred);
    }
    const int q = (int)rows.size();
    // wraps of ekf.cpp:304-307 hit state indices 2, 4, 6, ...: theta and every bearing row
    // (index 0, the x innovation, is NOT wrapped: the loop writes Y(2 + j), j >= 0)
    for (int k = 0; k < q; ++k)
        if (rows[k] >= 2 && rows[k] % 2 == 0) yv[k] = normalize_angle(c, (float)yv[k]);

    // HP = H_obs P (q x n), PHt = P H_obs^T (n x q)
    std::vector<double> HP((size_t)q * N), PHt((size_t)q * N);
#pragma omp parallel for
    for (int64_t j = 0; j < (int64_t)N; ++j)
        for (int k = 0; k < q; ++k) {
            double s = 0.0, t = 0.0;
            for (int e = 0; e < H[k].nnz; ++e) {
                s += H[k].v[e] * Pm(H[k].col[e], j);
                t += Pm(j, H[k].col[e]) * H[k].v[e];
            }
            HP[(size_t)k * N + j] = s;
            PHt[(size_t)k * N + j] = t;
        }
    Mat S(q, q);
    for (int a = 0; a < q; ++a)
        for (int b = 0; b < q; ++b) {
            double s = 0.0;
            for (int e = 0; e < H[b].nnz; ++e) s += HP[(size_t)a * N + H[b].col[e]] * H[b].v[e];
            S(a, b) = s + (a == b ? rdiag[a] : 0.0);
        }
    Mat Si = inverse(S);
    // K = PHt * Si (n x q), stored [k][i]
    std::vector<double> K((size_t)q * N);
#pragma omp parallel for
    for (int64_t i = 0; i < (int64_t)N; ++i)
        for (int k = 0; k < q; ++k) {
            double s = 0.0;
            for (int e = 0; e < q; ++e) s += PHt[(size_t)e * N + i] * Si(e, k);
            K[(size_t)k * N + i] = s;
        }
    for (size_t i = 0; i < N; ++i) {
        double s = 0.0;
        for (int k = 0; k < q; ++k) s += K[(size_t)k * N + i] * yv[k];
        X[i] += s;
    }
#pragma omp parallel for
    for (int64_t j = 0; j < (int64_t)N; ++j)
        for (size_t i = 0; i < N; ++i) {
            double s = 0.0;
            for (int k = 0; k < q; ++k) s += K[(size_t)k * N + i] * HP[(size_t)k * N + j];
            Pm(i, j) -= s;
        }
    return q;
}

// Laser front end (laser_front.hpp) over `nscan` scans of 360 ranges, OpenMP over scans. out_range / out_bearing:
// [nscan][cap]; counts [nscan] (-1: the reference's assert fires). Returns the wall time in seconds.
double orc_laser_extract(const float* ranges, int nscan, int cap, float* out_range, float* out_bearing, int* counts) {
    const LaserConfig cfg;
    laser_extract(cfg, ranges, out_range, out_bearing, 0);  // builds the sin / cos tables before the parallel loop
    const auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(static)
    for (int s = 0; s < nscan; ++s)
        counts[s] = laser_extract(cfg, ranges + (size_t)s * 360, out_range + (size_t)s * cap, out_bearing + (size_t)s * cap, cap);
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}  // extern "C"
