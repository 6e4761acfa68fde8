// aslam_filters.hpp -- C++ host-side mirror of the reference's filter classes over the C ABI (aslam_c.h).
//
// The reference exposes its filters as two ROS node classes with private callbacks:
//     aslam::EKFSlam   awesome_slam/src/ekf/ekf.h:71-132     cbSensorLandmark ekf.cpp:102-114, cbOdom ekf.cpp:74-99
//     aslam::UKFSlam   awesome_slam/src/ukf/ukf.h:86-144     cbSensorLandmark ukf.cpp:98-110,  cbOdom ukf.cpp:70-95
// and its plain types in include/awesome_slam/structures.h:44-112 (Point, LaserData) and
// awesome_slam_msgs/msg/Landmarks.msg:1-2. The classes below keep those names, the callback names (made public) and
// the argument meaning, with ROS message pointers replaced by plain structs holding exactly the fields the reference
// reads (ekf.cpp:94,139-142,206-209). Everything numerical happens in libaslam_b200.so on the GPU; this header only
// marshals. Header-only, C++14, no dependency beyond aslam_c.h. Errors: the reference has none (no exceptions, no
// return codes); here a failed C call throws aslam::Error carrying the aslam_status and aslam_last_error().
#ifndef ASLAM_FILTERS_HPP
#define ASLAM_FILTERS_HPP

#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "aslam_c.h"

namespace aslam {

struct Error : std::runtime_error {
    aslam_status status;
    Error(aslam_status s, const std::string& what) : std::runtime_error(what), status(s) {}
};
inline void check(aslam_status s) {
    if (s != ASLAM_OK) throw Error(s, std::string("aslam status ") + std::to_string(s) + ": " + aslam_last_error());
}

// structures.h:85-112 (range, bearing are floats there as well)
struct LaserData {
    float range = 0.f, bearing = 0.f;
    LaserData() {}
    LaserData(const float& a, const float& b) : range(a), bearing(b) {}
    void assign(const float& a, const float& b) {
        range = a;
        bearing = b;
    }
};

// awesome_slam_msgs/Landmarks: float64[] x, float64[] y. On out/landmarks/sensor x = range, y = bearing
// (sensor_landmark.cpp:127-132); on out/landmarks/kalman x, y are world coordinates (common.h:93-108).
struct Landmarks {
    std::vector<double> x, y;
};

// nav_msgs/Odometry, the fields the filters read
struct Odometry {
    struct {
        double x = 0, y = 0;
    } position;
    struct {
        double w = 1, x = 0, y = 0, z = 0;
    } orientation;
    double linear_x = 0;   // twist.twist.linear.x
    double angular_z = 0;  // twist.twist.angular.z
};

namespace detail {

class SingleFilter {
  public:
    // The reference's constructor takes nothing (constants come from config.h); `cfg` lifts them to run time.
    SingleFilter(int kind, const aslam_config* cfg, int device) {
        aslam_config c;
        if (cfg)
            c = *cfg;
        else
            aslam_config_default(&c);
        check(kind == ASLAM_KIND_EKF ? aslam_ekf_create(&c, 1, device, &h_) : aslam_ukf_create(&c, 1, device, &h_));
        ld_ = aslam_filter_ld(h_);
    }
    ~SingleFilter() {
        if (h_) aslam_filter_destroy(h_);
    }
    SingleFilter(const SingleFilter&) = delete;
    SingleFilter& operator=(const SingleFilter&) = delete;

    // cbSensorLandmark(msg): msg.x = ranges, msg.y = bearings; stored as float (ekf.cpp:111)
    void cbSensorLandmark(const Landmarks& msg) {
        const uint32_t n = (uint32_t)std::min(msg.x.size(), msg.y.size());
        r_.resize(n);
        b_.resize(n);
        for (uint32_t i = 0; i < n; ++i) {
            r_[i] = (float)msg.x[i];
            b_[i] = (float)msg.y[i];
        }
        const uint32_t off[2] = {0, n};
        check(aslam_filter_set_observations(h_, r_.data(), b_.data(), off));
        check(aslam_filter_sync(h_));
    }

    // cbOdom(msg): delta_time replaces the wall-clock difference of ekf.cpp:79-82 (the reference caps it at 1).
    // Returns false while gated by `if (init_z) return;` (ekf.cpp:76); otherwise fills `out` with the message the
    // reference publishes on out/landmarks/kalman (ekf.cpp:97-98).
    bool cbOdom(const Odometry& msg, float delta_time, Landmarks* out = nullptr) {
        const double x = msg.position.x, y = msg.position.y, vx = msg.linear_x, az = msg.angular_z;
        const float yaw = aslam_quat2euler(msg.orientation.w, msg.orientation.x, msg.orientation.y, msg.orientation.z);
        const float dt = std::min(delta_time, 1.0f);
        aslam_odom od{&x, &y, &yaw, &vx, &az, &dt};
        const aslam_status s = aslam_filter_step(h_, &od);
        if (s == ASLAM_ERR_NOT_READY) return false;
        check(s);
        if (out) {
            const uint32_t lmax = (ld_ - 3) / 2;
            out->x.assign(lmax, 0.0);
            out->y.assign(lmax, 0.0);
            uint32_t count = 0;
            check(aslam_filter_get_landmarks(h_, out->x.data(), out->y.data(), &count));
            out->x.resize(count);
            out->y.resize(count);
        } else {
            check(aslam_filter_sync(h_));
        }
        return true;
    }

    // param.X, param.P (column-major N x N, as Eigen stores it), N, and the per-observation corr_id of the last message
    struct State {
        uint32_t N = 0, status = 0;
        std::vector<double> X, P;
        std::vector<int32_t> assoc;
    };
    State state() {
        State s;
        std::vector<double> X(ld_), P((size_t)ld_ * ld_);
        s.assoc.assign(std::max<size_t>(r_.size(), 1), -2);
        check(aslam_filter_get_state(h_, X.data(), P.data(), &s.N, s.assoc.data(), &s.status));
        s.assoc.resize(r_.size());
        s.X.assign(X.begin(), X.begin() + s.N);
        s.P.resize((size_t)s.N * s.N);
        for (uint32_t j = 0; j < s.N; ++j)
            for (uint32_t i = 0; i < s.N; ++i) s.P[(size_t)j * s.N + i] = P[(size_t)j * ld_ + i];
        return s;
    }
    aslam_filter* handle() { return h_; }

  protected:
    aslam_filter* h_ = nullptr;
    uint32_t ld_ = 0;
    std::vector<float> r_, b_;
};

}  // namespace detail

class EKFSlam : public detail::SingleFilter {
  public:
    explicit EKFSlam(const aslam_config* cfg = nullptr, int device = 0) : SingleFilter(ASLAM_KIND_EKF, cfg, device) {}
};

class UKFSlam : public detail::SingleFilter {
  public:
    explicit UKFSlam(const aslam_config* cfg = nullptr, int device = 0) : SingleFilter(ASLAM_KIND_UKF, cfg, device) {}
    // first non-positive pivot of Paug.llt() in the last message, -1 if it factorised (ukf.cpp:280)
    int32_t lltFail() {
        int32_t k = -1;
        check(aslam_filter_get_llt_fail(h_, &k));
        return k;
    }
};

}  // namespace aslam
#endif  // ASLAM_FILTERS_HPP
