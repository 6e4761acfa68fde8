// Deterministic synthetic replay of SURVEY.md section 8d: the ROS-free stand-in for Gazebo's
// /odom (external/turtlebot3_simulations/.../turtlebot3_fake.cpp:161-177 lists the fields) and
// for the sensor_landmark node's (range, bearing) list (src/sensor_landmark/
// sensor_landmark.cpp:127-132). Host-only code; the same bytes feed the CUDA path, the
// oracle and the CPU baseline.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <random>
#include <thread>
#include <vector>

#include "../../include/aslam_c.h"

namespace {

struct Gauss {
    std::mt19937_64 rng;
    bool have = false;
    double spare = 0.0;
    explicit Gauss(uint64_t seed) : rng(seed) {}
    double uniform() { return ((double)(rng() >> 11) + 0.5) * (1.0 / 9007199254740992.0); }
    double operator()() {  // Box-Muller, both branches used
        if (have) {
            have = false;
            return spare;
        }
        const double u1 = uniform(), u2 = uniform();
        const double r = std::sqrt(-2.0 * std::log(u1));
        const double a = 6.283185307179586476925286766559 * u2;
        spare = r * std::sin(a);
        have = true;
        return r * std::cos(a);
    }
};

double wrap_pi(double a) {
    const double two_pi = 6.283185307179586476925286766559;
    a = std::fmod(a, two_pi);
    if (a > 3.14159265358979323846) a -= two_pi;
    if (a <= -3.14159265358979323846) a += two_pi;
    return a;
}

void one_filter(const double* lm_x, const double* lm_y, uint32_t L, uint32_t batch, uint32_t steps, uint32_t warm,
                double vx, double az, uint64_t seed, uint32_t b, double* ox, double* oy, float* yaw, double* ovx,
                double* oaz, float* dt, float* range, float* bearing, double* oqw, double* oqz) {
    Gauss g(seed);
    double x = 0.0, y = 0.0, th = 0.0;
    std::vector<std::pair<double, uint32_t>> order(L);
    for (uint32_t t = 0; t < steps; ++t) {
        const double v = t < warm ? 0.0 : vx, w = t < warm ? 0.0 : az;
        const double d = 1.0;
        if (w != 0.0) {  // exact unicycle arc over the interval that ends at this message
            const double th1 = th + w * d;
            x += v / w * (std::sin(th1) - std::sin(th));
            y -= v / w * (std::cos(th1) - std::cos(th));
            th = th1;
        } else {
            x += v * d * std::cos(th);
            y += v * d * std::sin(th);
        }
        const size_t i = (size_t)t * batch + b;
        ox[i] = x + 0.01 * g();
        oy[i] = y + 0.01 * g();
        const double nyaw = th + 0.01 * g();
        const double qw = std::cos(0.5 * nyaw), qz = std::sin(0.5 * nyaw);  // orientation (w, 0, 0, z) of the message
        yaw[i] = aslam_quat2euler(qw, 0.0, 0.0, qz);
        if (oqw) oqw[i] = qw;
        if (oqz) oqz[i] = qz;
        ovx[i] = v;
        oaz[i] = w;
        dt[i] = 1.0f;
        for (uint32_t l = 0; l < L; ++l)
            order[l] = {wrap_pi(std::atan2(lm_y[l] - y, lm_x[l] - x) - th), l};
        std::sort(order.begin(), order.end());
        for (uint32_t k = 0; k < L; ++k) {
            const uint32_t l = order[k].second;
            const double dx = lm_x[l] - x, dy = lm_y[l] - y;
            const double r = std::sqrt(dx * dx + dy * dy) + 0.01 * g();
            const double be = wrap_pi(order[k].first + 0.005 * g());
            range[i * L + k] = (float)r;     // ekf.cpp:111: the message doubles are stored as float
            bearing[i * L + k] = (float)be;
        }
    }
}

}  // namespace

// tools.h:62-66 quat2euler: the four arguments are rounded to float first, float arithmetic, atan2f.
extern "C" float aslam_quat2euler(double w, double x, double y, double z) {
    const float fw = (float)w, fx = (float)x, fy = (float)y, fz = (float)z;
    const float num = 2 * (fw * fz + fx * fy);
    const float den = 1 - 2 * (fz * fz + fy * fy);
    return std::atan2(num, den);
}

extern "C" aslam_status aslam_replay_generate_quat(const double* lm_x, const double* lm_y, uint32_t L, uint32_t batch,
                                                   uint32_t steps, uint32_t warm, double vx, double az,
                                                   uint64_t seed0, uint32_t first_filter, double* ox, double* oy,
                                                   float* yaw, double* ovx, double* oaz, float* dt, float* range,
                                                   float* bearing, double* qw, double* qz) {
    if (!ox || !oy || !yaw || !ovx || !oaz || !dt || (L && (!lm_x || !lm_y || !range || !bearing)))
        return ASLAM_ERR_ARG;
    unsigned nt = std::thread::hardware_concurrency();
    if (nt == 0) nt = 1;
    nt = std::min<unsigned>(nt, std::max<uint32_t>(1, batch / 64));
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < nt; ++t)
        pool.emplace_back([=]() {
            for (uint32_t b = t; b < batch; b += nt)
                one_filter(lm_x, lm_y, L, batch, steps, warm, vx, az, seed0 + first_filter + b, b, ox, oy, yaw, ovx,
                           oaz, dt, range, bearing, qw, qz);
        });
    for (auto& th : pool) th.join();
    return ASLAM_OK;
}

extern "C" aslam_status aslam_replay_generate(const double* lm_x, const double* lm_y, uint32_t L, uint32_t batch,
                                              uint32_t steps, uint32_t warm, double vx, double az, uint64_t seed0,
                                              uint32_t first_filter, double* ox, double* oy, float* yaw,
                                              double* ovx, double* oaz, float* dt, float* range, float* bearing) {
    return aslam_replay_generate_quat(lm_x, lm_y, L, batch, steps, warm, vx, az, seed0, first_filter, ox, oy, yaw,
                                      ovx, oaz, dt, range, bearing, nullptr, nullptr);
}
