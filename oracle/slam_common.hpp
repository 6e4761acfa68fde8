// TEST INFRASTRUCTURE ONLY -- part of the CPU oracle (see oracle/README.md).
//
// Restatement of the reference's shared scalar/model helpers. Every FP32 truncation site of
// SURVEY.md section 8a-casts is kept, in order. Reference paths are relative to
// /root/reference/awesome_slam/.
#pragma once
#include <cmath>
#include <cstdint>
#include <utility>
#include <vector>

#include "dense.hpp"

namespace orc {

// include/awesome_slam/config.h:39-65 -- every tunable is a `const float`.
struct Config {
    float pi = 3.141592654f;             // config.h:39
    float min_dist_thresh = 0.5f;        // config.h:43
    int min_landmark_occ = 10;           // config.h:44
    int max_state_dim = 30;              // config.h:45 MAX_LANDMARK_COUNT (compared to the STATE dim)
    float ukf_std_a = 0.2f;              // config.h:54
    float ukf_std_yaw = 0.2f;            // config.h:55
    float ukf_kp_robot_pose = 0.001f;    // config.h:56
    float ukf_kp_landmark_pose = 1.0f;   // config.h:57
    float ukf_kr = 0.2f;                 // config.h:58
    float ukf_kq = 0.001f;               // config.h:59
    float ekf_kp_robot_pose = 0.001f;    // config.h:62
    float ekf_kp_landmark_pose = 10000.0f;  // config.h:63 (dead: only scales a 0x0 block)
    float ekf_kr = 0.2f;                 // config.h:64
    float ekf_kq = 0.001f;               // config.h:65
};

// tools.h:44-50. Argument and result are float; fmodf; 2*PI is a float product.
inline float normalize_angle(const Config& cfg, float theta) {
    const float two_pi = 2 * cfg.pi;
    float ret = std::fmod(theta, two_pi);
    ret = ret > cfg.pi ? ret - two_pi : ret;
    ret = ret < -cfg.pi ? ret + two_pi : ret;
    return ret;
}

// structures.h:44-83 Point: a Vector2d whose only setters take `const float&`.
struct Pt {
    double x = 0, y = 0;
    Pt() {}
    Pt(float a, float b) : x(a), y(b) {}
};

// tools.h:53-59 (via structures.h:71-75): float dx, dy; sqrtf(dx*dx + dy*dy), no FMA.
inline float euler_distance(const Pt& p1, const Pt& p2) {
    float dx = (float)(p1.x - p2.x);
    float dy = (float)(p1.y - p2.y);
    float ret = std::sqrt(dx * dx + dy * dy);
    return ret;
}

// structures.h:85-112 LaserData
struct Laser {
    float range = 0, bearing = 0;
};

// structures.h:104-111: double evaluation, float store.
inline Pt to_point(const Laser& l, const Mat& Z) {
    float a = (float)(Z(0) + (double)l.range * std::cos(Z(2) + (double)l.bearing));
    float b = (float)(Z(1) + (double)l.range * std::sin(Z(2) + (double)l.bearing));
    return Pt(a, b);
}

// tools.h:62-66: all four components are rounded to float first, atan2f.
inline float quat2euler(float w, float x, float y, float z) {
    float yaw = std::atan2(2 * (w * z + x * y), 1 - 2 * (z * z + y * y));
    return yaw;
}

// common.h:46-75 stateTransitionFunction. vx, az, dt are floats; `point` may be augmented
// (size N+2). Only rows 0-2 change.
inline Mat state_transition(uint32_t N, const Mat& point, float vx, float az, float dt) {
    Mat P((int)N, 1);
    for (uint32_t i = 0; i < N; ++i) P(i) = point(i);
    if (std::fabs(az) > 0.001) {                      // common.h:52 (float |az| promoted)
        float r = vx / az;                            // common.h:54 f32 divide
        const float azdt = az * dt;                   // common.h:55-56 f32 multiply
        P(0) += (double)r * (-std::sin(point(2)) + std::sin(point(2) + (double)azdt));
        P(1) += (double)r * (std::cos(point(2)) - std::cos(point(2) + (double)azdt));
    } else {
        const float vxdt = vx * dt;                   // common.h:60-61 f32 multiply first
        P(0) += (double)vxdt * std::cos(point(2));
        P(1) += (double)vxdt * std::sin(point(2));
    }
    P(2) += (double)(float)(az * dt);                 // common.h:64
    if ((uint32_t)point.size() > N) {                 // common.h:67-72: all-double products
        P(0) += 0.5 * dt * dt * point(N) * std::cos(point(2));
        P(1) += 0.5 * dt * dt * point(N) * std::sin(point(2));
        P(2) += 0.5 * dt * dt * az;                   // uses az, not point(N+1), as written
    }
    return P;
}

// common.h:78-90 measurementFunction: identity on 0..2, range / un-wrapped bearing after.
inline Mat measurement(uint32_t N, const Mat& point) {
    Mat P = point;
    for (uint32_t i = 0; i < N - 3; i += 2) {
        P(3 + i) = std::sqrt(std::pow(point(3 + i) - point(0), 2) + std::pow(point(4 + i) - point(1), 2));
        P(4 + i) = std::atan2(point(4 + i) - point(1), point(3 + i) - point(0)) - point(2);
    }
    return P;
}

// Shared front half of both filters: ekf.cpp:137-203 / ukf.cpp:113-179 (association),
// ekf.cpp:217-253 / ukf.cpp:184-220 (wait-list). `grow` is the filter-specific
// updateNewLandmark. Records the per-observation decision for bit-exact comparison:
//   assoc[i] = j  (>= 0)  observation i was written into landmark slot j/2
//   assoc[i] = -1         sent to the wait-list
struct Frontend {
    Config cfg;
    uint32_t N = 3;
    std::vector<Laser> sensor_landmark;
    std::vector<std::pair<Laser, uint32_t>> wait;
    std::vector<int32_t> assoc;
    bool cap_rejected = false;

    void update_wait(const Laser& data, const Mat& Z) {
        if (wait.empty()) {
            wait.push_back({data, 1});
            return;
        }
        uint32_t corr = 0;
        Pt l2 = to_point(data, Z);
        Pt l1 = to_point(wait[0].first, Z);
        float mindist = euler_distance(l2, l1);
        const uint32_t size = (uint32_t)wait.size();
        for (uint32_t i = 1; i < size; ++i) {
            l1 = to_point(wait[i].first, Z);
            float dist = euler_distance(l2, l1);
            if (dist < mindist) {
                corr = i;
                mindist = dist;
            }
        }
        if (mindist < cfg.min_dist_thresh)
            wait[corr].second++;
        else
            wait.push_back({data, 1});
    }

    // Returns the list of landmarks promoted this step (may be empty).
    std::vector<Laser> associate(const Mat& X, Mat& Z) {
        assoc.assign(sensor_landmark.size(), -1);
        size_t oi = 0;
        for (Laser& data : sensor_landmark) {
            const size_t me = oi++;
            data.bearing = normalize_angle(cfg, data.bearing);
            if (N == 3) {
                update_wait(data, Z);
                continue;
            }
            int corr = 0;
            Pt l1((float)X(3), (float)X(4));
            Pt l2 = to_point(data, Z);
            float mindist = euler_distance(l2, l1);
            for (uint32_t j = 2; j < N - 3; j += 2) {
                l1 = Pt((float)X(3 + j), (float)X(4 + j));
                float dist = euler_distance(l2, l1);
                if (dist < mindist) {
                    corr = (int)j;
                    mindist = dist;
                }
            }
            if (mindist < cfg.min_dist_thresh) {
                Z(3 + corr) = data.range;
                Z(4 + corr) = data.bearing;
                assoc[me] = corr;
                continue;
            } else {
                update_wait(data, Z);
            }
        }
        std::vector<Laser> fresh;
        for (auto& w : wait) {
            if (w.second == (uint32_t)cfg.min_landmark_occ) {
                fresh.push_back(w.first);
                w.second += 1;
            }
        }
        return fresh;
    }
};

}  // namespace orc
