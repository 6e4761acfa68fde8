// aslam_replay -- ROS-free C++ harness: replays synthetic odometry + landmark observations
// through the filters of libaslam_b200.so. It stands where roscore + Gazebo + sensor_landmark + `rosrun awesome_slam
// ekf|ukf` stand in the reference (awesome_slam/launch/awesome_slam.launch, README.md:29-45): the map comes from the
// same landmarks.yaml, /odom and out/landmarks/sensor are generated, out/landmarks/kalman is printed / written as CSV.
//
//   aslam_replay [--filter ekf|ukf] [--map landmarks.yaml] [--steps 1000] [--batch 1] [--seed 1] [--cap 30]
//                [--device 0] [--csv out.csv]
//   --batch 1   message by message through aslam::EKFSlam / UKFSlam (cbSensorLandmark, cbOdom), like the ROS node
//   --batch B   B independent filters (per-filter noise seeds) in one launch (aslam_filter_run_host)
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "aslam_filters.hpp"

namespace {

// The two ROS-param lists of config/landmarks.yaml (block style "- v" or flow style "[a, b]").
bool load_map(const std::string& path, std::vector<double>& x, std::vector<double>& y) {
    std::ifstream f(path);
    if (!f) return false;
    std::string line;
    std::vector<double>* cur = nullptr;
    auto numbers = [](const std::string& s, std::vector<double>& out) {
        const char* p = s.c_str();
        while (*p) {
            char* e = nullptr;
            const double v = std::strtod(p, &e);
            if (e != p) {
                out.push_back(v);
                p = e;
            } else {
                ++p;
            }
        }
    };
    while (std::getline(f, line)) {
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.resize(hash);
        const size_t kx = line.find("landmarks/x"), ky = line.find("landmarks/y");
        if (kx != std::string::npos || ky != std::string::npos) {
            cur = kx != std::string::npos ? &x : &y;
            const size_t colon = line.find(':');
            if (colon != std::string::npos) numbers(line.substr(colon + 1), *cur);
        } else if (cur) {
            numbers(line, *cur);
        }
    }
    return !x.empty() && x.size() == y.size();
}

const char* arg(int argc, char** argv, const char* name, const char* def) {
    for (int i = 1; i + 1 < argc; ++i)
        if (!std::strcmp(argv[i], name)) return argv[i + 1];
    return def;
}

}  // namespace

int main(int argc, char** argv) {
    const std::string kind = arg(argc, argv, "--filter", "ekf");
    const std::string map = arg(argc, argv, "--map", "awesomeslam_b200/config/landmarks.yaml");
    const uint32_t steps = (uint32_t)std::atoi(arg(argc, argv, "--steps", "1000"));
    const uint32_t batch = (uint32_t)std::atoi(arg(argc, argv, "--batch", "1"));
    const uint64_t seed = (uint64_t)std::atoll(arg(argc, argv, "--seed", "1"));
    const int device = std::atoi(arg(argc, argv, "--device", "0"));
    const std::string csv = arg(argc, argv, "--csv", "");
    std::vector<double> lx, ly;
    if (!load_map(map, lx, ly)) {
        std::fprintf(stderr, "cannot read the landmark map %s\n", map.c_str());
        return 2;
    }
    const uint32_t L = (uint32_t)lx.size();
    aslam_config cfg;
    aslam_config_default(&cfg);
    cfg.max_state_dim = std::atoi(arg(argc, argv, "--cap", "30"));  // MAX_LANDMARK_COUNT (config.h:45)

    // synthetic /odom and out/landmarks/sensor: 15 stationary messages, then vx = 0.05, az = 0.1, dt = 1
    const size_t SB = (size_t)steps * batch;
    std::vector<double> ox(SB), oy(SB), vx(SB), az(SB);
    std::vector<float> yaw(SB), dt(SB), rng(SB * L), brg(SB * L);
    aslam::check(aslam_replay_generate(lx.data(), ly.data(), L, batch, steps, 15, 0.05, 0.1, seed, 0, ox.data(), oy.data(),
                                       yaw.data(), vx.data(), az.data(), dt.data(), rng.data(), brg.data()));
    std::vector<double> est_x, est_y;
    uint32_t N = 0, status = 0;
    const auto t0 = std::chrono::steady_clock::now();
    try {
        if (batch == 1) {
            aslam::EKFSlam* ekf = kind == "ekf" ? new aslam::EKFSlam(&cfg, device) : nullptr;
            aslam::UKFSlam* ukf = kind == "ukf" ? new aslam::UKFSlam(&cfg, device) : nullptr;
            if (!ekf && !ukf) {
                std::fprintf(stderr, "--filter must be ekf or ukf\n");
                return 2;
            }
            aslam::Landmarks sensed, out;
            std::ofstream cf;
            if (!csv.empty()) cf.open(csv);
            for (uint32_t t = 0; t < steps; ++t) {
                sensed.x.assign(rng.begin() + (size_t)t * L, rng.begin() + (size_t)(t + 1) * L);
                sensed.y.assign(brg.begin() + (size_t)t * L, brg.begin() + (size_t)(t + 1) * L);
                aslam::Odometry od;
                od.position.x = ox[t];
                od.position.y = oy[t];
                od.orientation.w = std::cos(0.5 * (double)yaw[t]);  // yaw-only quaternion
                od.orientation.z = std::sin(0.5 * (double)yaw[t]);
                od.linear_x = vx[t];
                od.angular_z = az[t];
                if (ekf) {
                    ekf->cbSensorLandmark(sensed);
                    ekf->cbOdom(od, dt[t], &out);
                } else {
                    ukf->cbSensorLandmark(sensed);
                    ukf->cbOdom(od, dt[t], &out);
                }
                if (cf.is_open()) {
                    cf << t;
                    for (size_t i = 0; i < out.x.size(); ++i) cf << ',' << out.x[i] << ',' << out.y[i];
                    cf << '\n';
                }
            }
            auto st = ekf ? ekf->state() : ukf->state();
            N = st.N;
            status = st.status;
            est_x = out.x;
            est_y = out.y;
            delete ekf;
            delete ukf;
        } else {
            aslam_filter* h = nullptr;
            aslam::check(kind == "ekf" ? aslam_ekf_create(&cfg, batch, device, &h) : aslam_ukf_create(&cfg, batch, device, &h));
            aslam_odom od{ox.data(), oy.data(), yaw.data(), vx.data(), az.data(), dt.data()};
            aslam::check(aslam_filter_run_host(h, steps, &od, rng.data(), brg.data(), L));
            const uint32_t ld = aslam_filter_ld(h), lmax = (ld - 3) / 2;
            std::vector<double> x((size_t)batch * lmax), y((size_t)batch * lmax);
            std::vector<uint32_t> cnt(batch), Ns(batch), sts(batch);
            aslam::check(aslam_filter_get_landmarks(h, x.data(), y.data(), cnt.data()));
            aslam::check(aslam_filter_get_state(h, nullptr, nullptr, Ns.data(), nullptr, sts.data()));
            est_x.assign(x.begin(), x.begin() + cnt[0]);
            est_y.assign(y.begin(), y.begin() + cnt[0]);
            N = Ns[0];
            for (uint32_t b = 0; b < batch; ++b) status |= sts[b];
            aslam_filter_destroy(h);
        }
    } catch (const aslam::Error& e) {
        std::fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::printf("%s-slam: %u filter(s) x %u messages in %.3f s (%.0f updates/s), state dim N = %u, status 0x%x\n",
                kind.c_str(), batch, steps, sec, (double)batch * steps / sec, N, status);
    // every estimate against its nearest map landmark (what visualize_landmark shows as green vs blue cylinders)
    double worst = 0.0;
    for (size_t i = 0; i < est_x.size(); ++i) {
        double best = 1e300;
        for (uint32_t l = 0; l < L; ++l) best = std::min(best, std::hypot(est_x[i] - lx[l], est_y[i] - ly[l]));
        std::printf("  landmark %2zu  (% .4f, % .4f)  error %.4f m\n", i, est_x[i], est_y[i], best);
        worst = std::max(worst, best);
    }
    std::printf("mapped %zu of %u landmarks, worst error %.4f m\n", est_x.size(), L, worst);
    return 0;
}
