// TEST INFRASTRUCTURE ONLY -- part of the CPU oracle (see oracle/README.md).
//
// ROS-free, Eigen-free restatement of aslam::EKFSlam (src/ekf/ekf.cpp:49-311) and
// aslam::UKFSlam (src/ukf/ukf.cpp:49-392, src/ukf/ukf.h:56-82). Callbacks take POD inputs:
//   set_observations  <-> cbSensorLandmark (ekf.cpp:102-114 / ukf.cpp:98-110)
//   step              <-> cbOdom           (ekf.cpp:74-99  / ukf.cpp:70-95)
// The yaw argument is the float produced by quat2euler (tools.h:62-66) on the host, and dt
// is the float delta_time of ekf.cpp:80 -- the replay supplies both directly.
// The only constant that differs from the reference is the run-time `max_state_dim`
// (reference: MAX_LANDMARK_COUNT = 30, config.h:45).
#pragma once
#include "slam_common.hpp"

namespace orc {

struct StepLog {
    int llt_fail_k = -1;  // UKF: first non-positive pivot column of Paug (-1: factorised)
};

class EkfOracle {
  public:
    Frontend fe;
    bool init_x = true, init_z = true;
    Mat X, Z, I, A, P, H, Q, R;

    explicit EkfOracle(const Config& cfg) {
        fe.cfg = cfg;
        initialize();
    }

    // ekf.cpp:49-71
    void initialize() {
        const Config& c = fe.cfg;
        fe.N = 3;
        init_x = init_z = true;
        X = Mat::zero(3, 1);
        Z = Mat::zero(3, 1);
        Q = Mat::zero(3, 3);
        A = Mat::identity(3, 3);
        H = Mat::identity(3, 3);
        I = Mat::identity(3, 3);
        R = scale(Mat::identity(3, 3), (double)c.ekf_kr);
        P = scale(Mat::identity(3, 3), (double)c.ekf_kp_robot_pose);
        Q(0, 0) = c.ekf_kq;
        Q(1, 1) = c.ekf_kq;
        Q(2, 2) = c.ekf_kq;
        // ekf.cpp:70 assigns a 0x0 block: no effect.
    }

    // ekf.cpp:102-114 (the f64 -> f32 cast of ekf.cpp:111 is done by the caller)
    void set_observations(const float* range, const float* bearing, uint32_t n) {
        init_z = false;
        fe.sensor_landmark.clear();
        for (uint32_t i = 0; i < n; ++i) fe.sensor_landmark.push_back(Laser{range[i], bearing[i]});
    }

    // ekf.cpp:117-134
    void update_h() {
        const uint32_t N = fe.N;
        for (uint32_t i = 0; i < N - 3; i += 2) {
            float hyp = (float)(std::pow(X(3 + i) - X(0), 2) + std::pow(X(4 + i) - X(1), 2));
            float dist = std::sqrt(hyp);
            H(3 + i, 0) = (-X(3 + i) + X(0)) / dist;
            H(4 + i, 0) = -(-X(4 + i) + X(1)) / hyp;
            H(3 + i, 1) = (-X(4 + i) + X(1)) / dist;
            H(4 + i, 1) = (-X(3 + i) + X(0)) / hyp;
            H(4 + i, 2) = -1;
            H(3 + i, 3 + i) = -(-X(3 + i) + X(0)) / dist;
            H(3 + i, 4 + i) = -(-X(4 + i) + X(1)) / dist;
            H(4 + i, 3 + i) = (-X(4 + i) + X(1)) / hyp;
            H(4 + i, 4 + i) = -(-X(3 + i) + X(0)) / hyp;
        }
    }

    // ekf.cpp:255-290
    void grow(const std::vector<Laser>& fresh) {
        const Config& c = fe.cfg;
        const uint32_t cacheN = fe.N;
        const uint32_t size = (uint32_t)fresh.size();
        fe.N += 2 * size;
        if (fe.N >= (uint32_t)c.max_state_dim) {
            fe.N = cacheN;
            fe.cap_rejected = true;
            return;
        }
        const int N = (int)fe.N;
        conservative_resize_like(X, Mat::zero(N, 1));
        conservative_resize_like(Z, Mat::zero(N, 1));
        conservative_resize_like(Q, Mat::zero(N, N));
        conservative_resize_like(I, Mat::identity(N, N));
        conservative_resize_like(A, Mat::identity(N, N));
        conservative_resize_like(H, Mat::identity(N, N));
        conservative_resize_like(P, scale(Mat::identity(N, N), (double)c.ukf_kp_landmark_pose));  // :277 (UKF const, as written)
        conservative_resize_like(R, scale(Mat::identity(N, N), (double)c.ukf_kr));                // :278
        for (uint32_t i = 0; i < size * 2; i += 2) {
            Z(cacheN + i) = fresh[i / 2].range;
            Z(cacheN + i + 1) = fresh[i / 2].bearing;
            X(cacheN + i) = Z(0) + Z(cacheN + i) * std::cos(Z(2) + Z(cacheN + i + 1));
            X(cacheN + i + 1) = Z(1) + Z(cacheN + i) * std::sin(Z(2) + Z(cacheN + i + 1));
        }
    }

    // ekf.cpp:137-213; vx, az are the message doubles here.
    void update_z_and_a(double ox, double oy, float yaw, double vx, double az, float dt) {
        Z(0) = ox;
        Z(1) = oy;
        Z(2) = yaw;
        std::vector<Laser> fresh = fe.associate(X, Z);
        if (fresh.size()) grow(fresh);
        if (vx && az) {
            float delta_theta = (float)(az * (double)dt);  // :208 f64*f32 -> f32
            float r = (float)(vx / az);                    // :209 f64/f64 -> f32
            A(0, 0) = (double)r * (-std::cos(Z(2)) + std::cos(Z(2) + (double)delta_theta));
            A(1, 0) = (double)r * (-std::sin(Z(2)) + std::sin(Z(2) + (double)delta_theta));
        }
    }

    // ekf.cpp:293-311; vx, az, dt arrive as floats (signature of slam()).
    void slam(float vx, float az, float dt) {
        const uint32_t N = fe.N;
        X = state_transition(N, X, vx, az, dt);
        X(2) = normalize_angle(fe.cfg, (float)X(2));
        P = add(matmul(matmul(A, P), transpose(A)), Q);
        update_h();
        Mat S = add(matmul(matmul(H, P), transpose(H)), R);
        Mat K = matmul(matmul(P, transpose(H)), inverse(S));
        Mat Y = sub(Z, measurement(N, X));
        for (uint32_t j = 0; j < N - 1; j += 2) Y(2 + j) = normalize_angle(fe.cfg, (float)Y(2 + j));
        X = add(X, matmul(K, Y));
        P = matmul(sub(I, matmul(K, H)), P);
    }

    // ekf.cpp:74-99. Returns false when gated by init_z (no observation message yet).
    bool step(double ox, double oy, float yaw, double vx, double az, float dt) {
        if (init_z) return false;
        update_z_and_a(ox, oy, yaw, vx, az, dt);
        if (init_x) {
            init_x = false;
            X = Z;
        }
        slam((float)vx, (float)az, dt);
        return true;
    }
};

class UkfOracle {
  public:
    Frontend fe;
    bool init_x = true, init_z = true;
    Mat X, Z, weights, P, Q, R;
    float lambda = 0;
    StepLog log;

    explicit UkfOracle(const Config& cfg) {
        fe.cfg = cfg;
        update_weights(3);  // ukf.h:68-71
        initialize();
    }

    // ukf.h:73-81: lambda, the common weight and w0 are all computed in float.
    void update_weights(uint32_t N) {
        lambda = (float)(3.0 - (double)(N + 2));
        float weight = (float)(0.5 / (double)(lambda + (float)N + 2));  // (lambda + N) + 2 in float
        weights = Mat::constant((int)(2 * N + 5), 1, 1.0);
        weights = scale(weights, (double)weight);
        weights(0) = lambda / (lambda + (float)N + 2);
    }

    // ukf.cpp:49-67
    void initialize() {
        const Config& c = fe.cfg;
        fe.N = 3;
        init_x = init_z = true;
        X = Mat::zero(3, 1);
        Z = Mat::zero(3, 1);
        Q = Mat::zero(3, 3);
        R = scale(Mat::identity(3, 3), (double)c.ukf_kr);
        P = scale(Mat::identity(3, 3), (double)c.ukf_kp_robot_pose);
        Q(0, 0) = c.ukf_kq;
        Q(1, 1) = c.ukf_kq;
        Q(2, 2) = c.ukf_kq;
    }

    void set_observations(const float* range, const float* bearing, uint32_t n) {
        init_z = false;
        fe.sensor_landmark.clear();
        for (uint32_t i = 0; i < n; ++i) fe.sensor_landmark.push_back(Laser{range[i], bearing[i]});
    }

    // ukf.cpp:222-257
    void grow(const std::vector<Laser>& fresh) {
        const Config& c = fe.cfg;
        const uint32_t cacheN = fe.N;
        const uint32_t size = (uint32_t)fresh.size();
        fe.N += 2 * size;
        if (fe.N >= (uint32_t)c.max_state_dim) {
            fe.N = cacheN;
            fe.cap_rejected = true;
            return;
        }
        const int N = (int)fe.N;
        conservative_resize_like(X, Mat::zero(N, 1));
        conservative_resize_like(Z, Mat::zero(N, 1));
        conservative_resize_like(Q, Mat::zero(N, N));
        conservative_resize_like(P, scale(Mat::identity(N, N), (double)c.ukf_kp_landmark_pose));
        conservative_resize_like(R, scale(Mat::identity(N, N), (double)c.ukf_kr));
        update_weights(fe.N);
        for (uint32_t i = 0; i < size * 2; i += 2) {
            Z(cacheN + i) = fresh[i / 2].range;
            Z(cacheN + i + 1) = fresh[i / 2].bearing;
            X(cacheN + i) = Z(0) + Z(cacheN + i) * std::cos(Z(2) + Z(cacheN + i + 1));
            X(cacheN + i + 1) = Z(1) + Z(cacheN + i) * std::sin(Z(2) + Z(cacheN + i + 1));
        }
    }

    // ukf.cpp:113-180
    void update_z(double ox, double oy, float yaw) {
        Z(0) = ox;
        Z(1) = oy;
        Z(2) = yaw;
        std::vector<Laser> fresh = fe.associate(X, Z);
        if (fresh.size()) grow(fresh);
    }

    void wrap_bearings(Mat& v) const {  // the `for j < N-1; j += 2: v(2+j) = normalizeAngle` idiom
        for (uint32_t j = 0; j < fe.N - 1; j += 2) v(2 + j) = normalize_angle(fe.cfg, (float)v(2 + j));
    }

    // ukf.cpp:260-392
    void slam(float vx, float az, float dt) {
        const Config& c = fe.cfg;
        const int N = (int)fe.N, NA = N + 2, NS = 2 * N + 5;
        Mat Xaug(NA, 1), Paug(NA, NA), XsigAug(NA, NS);
        for (int i = 0; i < N; ++i) Xaug(i) = X(i);
        Xaug(N) = 0.0;
        Xaug(N + 1) = 0.0;
        Paug.fill(0.0);
        for (int j = 0; j < N; ++j)
            for (int i = 0; i < N; ++i) Paug(i, j) = P(i, j);
        Paug(N, N) = c.ukf_std_a * c.ukf_std_a;          // :276 f32*f32
        Paug(N + 1, N + 1) = c.ukf_std_yaw * c.ukf_std_yaw;
        Mat L = llt_lower(Paug, &log.llt_fail_k);         // :280

        for (int i = 0; i < NA; ++i) XsigAug(i, 0) = Xaug(i);
        float w = std::sqrt(lambda + (float)(uint32_t)N + 2);  // :284 (float + uint32 -> float), sqrtf
        for (int i = 0; i < NA; ++i)
            for (int k = 0; k < NA; ++k) {
                XsigAug(k, i + 1) = Xaug(k) + (double)w * L(k, i);
                XsigAug(k, i + 3 + N) = Xaug(k) - (double)w * L(k, i);
            }

        Mat XsigPred(N, NS);                              // :292-297
        Mat col(NA, 1);
        for (int i = 0; i < NS; ++i) {
            for (int k = 0; k < NA; ++k) col(k) = XsigAug(k, i);
            Mat p = state_transition((uint32_t)N, col, vx, az, dt);
            for (int k = 0; k < N; ++k) XsigPred(k, i) = p(k);
            XsigPred(2, i) = normalize_angle(c, (float)XsigPred(2, i));
        }

        X.fill(0.0);                                      // :300-304
        for (int i = 0; i < NS; ++i)
            for (int k = 0; k < N; ++k) X(k) += weights(i) * XsigPred(k, i);

        Mat Xdiff(N, 1);                                  // :307-319
        P.fill(0.0);
        for (int i = 0; i < NS; ++i) {
            for (int k = 0; k < N; ++k) Xdiff(k) = XsigPred(k, i) - X(k);
            Xdiff(2) = normalize_angle(c, (float)Xdiff(2));
            for (int b = 0; b < N; ++b)
                for (int a = 0; a < N; ++a) P(a, b) += (weights(i) * Xdiff(a)) * Xdiff(b);
        }
        P = add(P, Q);

        Mat Zsig(N, NS);                                  // :322-326
        Mat pc(N, 1);
        for (int i = 0; i < NS; ++i) {
            for (int k = 0; k < N; ++k) pc(k) = XsigPred(k, i);
            Mat z = measurement((uint32_t)N, pc);
            for (int k = 0; k < N; ++k) Zsig(k, i) = z(k);
        }

        Mat Zpred = Mat::zero(N, 1);                      // :329-339
        for (int i = 0; i < NS; ++i)
            for (int k = 0; k < N; ++k) Zpred(k) += weights(i) * Zsig(k, i);
        wrap_bearings(Zpred);

        Mat Zdiff(N, 1);                                  // :342-357
        Mat S = Mat::zero(N, N);
        for (int i = 0; i < NS; ++i) {
            for (int k = 0; k < N; ++k) Zdiff(k) = Zsig(k, i) - Zpred(k);
            wrap_bearings(Zdiff);
            for (int b = 0; b < N; ++b)
                for (int a = 0; a < N; ++a) S(a, b) = S(a, b) + (weights(i) * Zdiff(a)) * Zdiff(b);
        }
        S = add(S, R);

        Mat Tc = Mat::zero(N, N);                         // :360-375
        for (int i = 0; i < NS; ++i) {
            for (int k = 0; k < N; ++k) Zdiff(k) = Zsig(k, i) - Zpred(k);
            wrap_bearings(Zdiff);
            for (int k = 0; k < N; ++k) Xdiff(k) = XsigPred(k, i) - X(k);
            Xdiff(2) = normalize_angle(c, (float)Xdiff(2));
            for (int b = 0; b < N; ++b)
                for (int a = 0; a < N; ++a) Tc(a, b) = Tc(a, b) + (weights(i) * Xdiff(a)) * Zdiff(b);
        }

        Mat K = matmul(Tc, inverse(S));                   // :378
        for (int k = 0; k < N; ++k) Zdiff(k) = Z(k) - Zpred(k);  // :381-386
        wrap_bearings(Zdiff);
        X = add(X, matmul(K, Zdiff));                     // :389
        P = sub(P, matmul(matmul(K, S), transpose(K)));   // :391
    }

    // ukf.cpp:70-95
    bool step(double ox, double oy, float yaw, double vx, double az, float dt) {
        if (init_z) return false;
        update_z(ox, oy, yaw);
        if (init_x) {
            init_x = false;
            X = Z;
        }
        slam((float)vx, (float)az, dt);
        return true;
    }
};

}  // namespace orc
