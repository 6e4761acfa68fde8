// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's laser front end:
// awesome_slam/src/sensor_landmark/sensor_landmark.cpp (LaserScan -> clusters -> circle test -> Chernov hyper-fit
// -> (range, bearing)), include/awesome_slam/structures.h:44-134 (Point, Circle) and tools.h:53-59. ROS-free and
// Eigen-free; every float/double rounding site of the reference is kept. Parity is UNPINNED by the reference (it
// ships no tests); the independent check is oracle/np_restatement.py::laser_extract (numpy SVD / eigh / solve).
//
// Eigen calls restated (Eigen3 is absent from this image, implied version 3.3.7):
//   JacobiSVD(Z) (sensor_landmark.cpp:251): singular values / V of the n x 4 data matrix, here from a cyclic Jacobi
//     eigen-decomposition of Z'Z (V, sigma = sqrt(lambda)); the perturbation this adds is O(eps |Z'Z|), far below
//     the float rounding of the published (range, bearing).
//   SelfAdjointEigenSolver(Q) (:272): cyclic Jacobi, eigenvalues sorted ascending as Eigen returns them.
//   Y.colPivHouseholderQr().solve(A_star) (:286): Y = V Sigma V' is symmetric positive definite, the solve is
//     V Sigma^-1 V' A_star.
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>

namespace orc {

struct LaserConfig {
    float deg2rad = 0.01745329251f;   // config.h:40
    float min_dist_thresh = 0.5f;     // config.h:43
    float std_max = 0.4f;             // config.h:48 STD
    float min_mean = 1.5f;            // config.h:49
    float max_mean = 3.0f;            // config.h:50
    int min_cluster_points = 3;       // config.h:51
};

struct LPoint {  // structures.h:44-83: Vector2d holding float-exact values
    double x, y;
};

// tools.h:53-59 on two Points
inline float laser_distance(const LPoint& p, const LPoint& q) {
    const float dx = (float)(p.x - q.x);
    const float dy = (float)(p.y - q.y);
    return std::sqrt(dx * dx + dy * dy);
}

// Cyclic Jacobi for a symmetric 4 x 4 matrix: A = V diag(w) V'. Eigenvalues unsorted.
inline void jacobi4(const double Ain[4][4], double w[4], double V[4][4]) {
    double A[4][4];
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) {
            A[i][j] = Ain[i][j];
            V[i][j] = i == j ? 1.0 : 0.0;
        }
    // Classical cyclic Jacobi: a rotation annihilates A[p][q] (set to exactly 0 afterwards); from the 4th sweep on an
    // off-diagonal entry that no longer changes either of its diagonal entries is dropped. Ends when the
    // off-diagonal part is exactly zero.
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0;
        for (int i = 0; i < 3; ++i)
            for (int j = i + 1; j < 4; ++j) off += std::fabs(A[i][j]);
        if (off == 0.0) break;
        for (int p = 0; p < 3; ++p)
            for (int q = p + 1; q < 4; ++q) {
                if (A[p][q] == 0.0) continue;
                const double g = 100.0 * std::fabs(A[p][q]);
                if (sweep > 3 && std::fabs(A[p][p]) + g == std::fabs(A[p][p]) &&
                    std::fabs(A[q][q]) + g == std::fabs(A[q][q])) {
                    A[p][q] = A[q][p] = 0.0;
                    continue;
                }
                const double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
                const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                const double c = 1.0 / std::sqrt(t * t + 1.0), s = t * c;
                for (int k = 0; k < 4; ++k) {  // A <- A J
                    const double akp = A[k][p], akq = A[k][q];
                    A[k][p] = c * akp - s * akq;
                    A[k][q] = s * akp + c * akq;
                }
                for (int k = 0; k < 4; ++k) {  // A <- J' A
                    const double apk = A[p][k], aqk = A[q][k];
                    A[p][k] = c * apk - s * aqk;
                    A[q][k] = s * apk + c * aqk;
                }
                A[p][q] = A[q][p] = 0.0;
                for (int k = 0; k < 4; ++k) {
                    const double vkp = V[k][p], vkq = V[k][q];
                    V[k][p] = c * vkp - s * vkq;
                    V[k][q] = s * vkp + c * vkq;
                }
            }
    }
    for (int i = 0; i < 4; ++i) w[i] = A[i][i];
}

// sensor_landmark.cpp:147-186 circleClassification (Xavier et al. 2005 inscribed-angle test), float arithmetic.
inline bool circle_classification(const LaserConfig& cfg, const std::vector<LPoint>& pts) {
    float sum = 0.0f;
    const int size = (int)pts.size();
    const LPoint P1 = pts.front(), P2 = pts.back();
    std::vector<float> angles;
    for (int i = 1; i < size - 1; ++i) {
        const LPoint& P = pts[i];
        const float c = laser_distance(P1, P2);
        const float a = laser_distance(P1, P);
        const float b = laser_distance(P2, P);
        const float angle = std::acos((a * a + b * b - c * c) / (2 * a * b));  // float overload (:164)
        angles.push_back(angle);
        sum += angle;
    }
    float variance = 0.0f;
    const float mean = sum / (size - 2);
    for (const float val : angles) variance += std::pow(val - mean, 2);  // pow(float, int) -> double (:174)
    const float sd = std::sqrt(variance / 5.0);                            // double sqrt, float store (:177)
    return sd < cfg.std_max && mean > cfg.min_mean && mean < cfg.max_mean;
}

// sensor_landmark.cpp:192-298 circleFitting (Al-Sharadqah & Chernov 2009 hyper-fit) followed by
// Circle::toLaserData (structures.h:126-133). Returns false when A(0) == 0 never happens in the reference
// (it would publish inf/nan): the values are returned as computed.
inline void circle_fitting(const std::vector<LPoint>& pts_in, float& out_range, float& out_bearing) {
    std::vector<LPoint> pts = pts_in;
    const size_t n = pts.size();
    float x_mean = 0.0f, y_mean = 0.0f, z_mean = 0.0f;
    for (const LPoint& p : pts) {
        x_mean += (float)p.x;  // float += float (Point::operator() returns float)
        y_mean += (float)p.y;
    }
    x_mean = x_mean / (float)n;
    y_mean = y_mean / (float)n;
    for (LPoint& p : pts) {  // pt.assign(pt(0) - x_mean, pt(1) - y_mean): float subtraction (:213)
        const float a = (float)p.x - x_mean, b = (float)p.y - y_mean;
        p.x = a;
        p.y = b;
    }
    std::vector<float> z;
    for (const LPoint& p : pts) {
        const float px = (float)p.x, py = (float)p.y;
        z.push_back(px * px + py * py);  // float arithmetic (:220)
    }
    for (const float zi : z) z_mean += zi;
    z_mean = z_mean / (float)z.size();
    // Z = [z x y 1] (n x 4); G = Z'Z
    double G[4][4] = {};
    for (size_t i = 0; i < n; ++i) {
        const double row[4] = {(double)z[i], pts[i].x, pts[i].y, 1.0};
        for (int a = 0; a < 4; ++a)
            for (int b = 0; b < 4; ++b) G[a][b] += row[a] * row[b];
    }
    // H^-1 (:245): [0 0 0 1/2; 0 1 0 0; 0 0 1 0; 1/2 0 0 -2 z_mean]
    double Hinv[4][4] = {};
    Hinv[0][3] = 0.5;
    Hinv[1][1] = 1.0;
    Hinv[2][2] = 1.0;
    Hinv[3][0] = 0.5;
    Hinv[3][3] = -2 * z_mean;  // int * float -> float, then widened
    double lam[4], V[4][4];
    jacobi4(G, lam, V);
    double sv[4];
    for (int i = 0; i < 4; ++i) sv[i] = std::sqrt(lam[i] > 0.0 ? lam[i] : 0.0);
    double sv_min = sv[0];
    int i_min = 0;
    for (int i = 1; i < 4; ++i)
        if (sv[i] < sv_min) {
            sv_min = sv[i];
            i_min = i;
        }
    double A[4];
    if (sv_min > 10e-12) {  // singularValues(3) (:256): the smallest, JacobiSVD sorts descending
        double Y[4][4] = {}, Q[4][4] = {}, T[4][4] = {};
        for (int a = 0; a < 4; ++a)
            for (int b = 0; b < 4; ++b)
                for (int k = 0; k < 4; ++k) Y[a][b] += V[a][k] * sv[k] * V[b][k];
        for (int a = 0; a < 4; ++a)
            for (int b = 0; b < 4; ++b)
                for (int k = 0; k < 4; ++k) T[a][b] += Y[a][k] * Hinv[k][b];
        for (int a = 0; a < 4; ++a)
            for (int b = 0; b < 4; ++b)
                for (int k = 0; k < 4; ++k) Q[a][b] += T[a][k] * Y[k][b];
        for (int a = 0; a < 4; ++a)  // SelfAdjointEigenSolver reads the lower triangle: make that explicit
            for (int b = a + 1; b < 4; ++b) Q[a][b] = Q[b][a];
        double ev[4], E[4][4];
        jacobi4(Q, ev, E);
        int order[4] = {0, 1, 2, 3};  // ascending, as Eigen returns them
        for (int a = 0; a < 4; ++a)
            for (int b = a + 1; b < 4; ++b)
                if (ev[order[b]] < ev[order[a]]) {
                    const int t = order[a];
                    order[a] = order[b];
                    order[b] = t;
                }
        float smallest_id = 0;  // (:259-260) both are floats in the reference
        float smallest_ev = 99999;
        for (int i = 0; i < 4; ++i)
            if (ev[order[i]] > 0 && ev[order[i]] < smallest_ev) {
                smallest_id = (float)i;
                smallest_ev = (float)ev[order[i]];
            }
        const int col = order[(int)smallest_id];
        // A = Y^-1 A_star = V Sigma^-1 V' A_star
        double t4[4];
        for (int k = 0; k < 4; ++k) {
            double d = 0.0;
            for (int a = 0; a < 4; ++a) d += V[a][k] * E[a][col];
            t4[k] = d / sv[k];
        }
        for (int a = 0; a < 4; ++a) {
            A[a] = 0.0;
            for (int k = 0; k < 4; ++k) A[a] += V[a][k] * t4[k];
        }
    } else {
        for (int a = 0; a < 4; ++a) A[a] = V[a][i_min];  // V.col(3)
    }
    const float a = (float)((-A[1]) / (2 * A[0]));
    const float b = (float)((-A[2]) / (2 * A[0]));
    const float cx = a + x_mean, cy = b + y_mean;  // Circle(const float&, const float&)
    const double dcx = cx, dcy = cy;
    out_range = (float)std::sqrt(dcx * dcx + dcy * dcy);  // structures.h:128
    out_bearing = (float)std::atan2(dcy, dcx);            // :129
}

// sensor_landmark.cpp:59-133 callback. ranges: 360 floats (inf = no return). cos_map / sin_map as initialize()
// builds them (:50-57). Returns the number of landmarks, or -1 where the reference's assert(theta < 359) fires
// (first and last ray within MIN_DIST_THRESH: its first-cluster loop then swallows the whole scan).
inline int laser_extract(const LaserConfig& cfg, const float* ranges, float* out_range, float* out_bearing, int cap) {
    static std::vector<float> sin_map, cos_map;
    if (sin_map.empty())
        for (uint16_t th = 0; th < 360; ++th) {
            sin_map.push_back(std::sin(cfg.deg2rad * static_cast<float>(th)));
            cos_map.push_back(std::cos(cfg.deg2rad * static_cast<float>(th)));
        }
    auto pose = [&](int th, int tab) {  // bearing2pose (:136-142) of ray th with table entry tab
        const float c = tab < 360 ? cos_map[tab] : 0.0f, s = tab < 360 ? sin_map[tab] : 0.0f;
        const float x = ranges[th] * c;
        const float y = ranges[th] * s;
        return LPoint{x, y};
    };
    // (:87,98) `bearing2pose(theta, msg->ranges[theta++])`: the evaluation order of the two arguments is
    // unspecified, and GCC -- the compiler of every ROS build and of oracle/_ref, which pins this -- takes them
    // right to left: ray t is projected with table entry t + 1 (a +1 degree rotation of every cluster and of the
    // published bearings), and ray 359 with entry 360, one past the table, read as 0.0f (fresh heap; oracle/_ref
    // sets it). The wrap test (:63-64) has no side effect in its arguments and uses entries 0 and 359.
    LPoint p1 = pose(0, 0), p2 = pose(359, 359);
    if (laser_distance(p1, p2) < cfg.min_dist_thresh) return -1;
    int theta = 0, count = 0;
    p1 = pose(theta, theta + 1);
    ++theta;
    std::vector<LPoint> cluster = {p1};
    while (theta < 360) {
        p2 = pose(theta, theta + 1);
        ++theta;
        if (laser_distance(p1, p2) < cfg.min_dist_thresh) {
            cluster.push_back(p2);
        } else if ((int)cluster.size() > cfg.min_cluster_points && circle_classification(cfg, cluster)) {
            // (theta == 360: joined with first_cluster, which is empty on this path)
            if (count < cap) circle_fitting(cluster, out_range[count], out_bearing[count]);
            ++count;
            cluster.clear();
        } else {
            cluster.clear();
        }
        p1 = p2;
    }
    return count;
}

}  // namespace orc
