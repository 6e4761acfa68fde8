// TEST INFRASTRUCTURE ONLY -- builds the reference's OWN laser front end (src/sensor_landmark/sensor_landmark.cpp
// with sensor_landmark.h, structures.h, tools.h, config.h, compiled from where they lie under /root/reference,
// unedited) into oracle/_ref/libref_laser.so, against the Eigen / ROS stand-ins of oracle/ref_shim/include
// (JacobiSVD / SelfAdjointEigenSolver / colPivHouseholderQr are restated in aslam_ref_shim/eigen_decomp.hpp).
#define main aslam_ref_laser_node_main  // sensor_landmark.cpp:300 defines main(); it is never called
#include <sensor_landmark/sensor_landmark.cpp>
#undef main

#include <chrono>
#include <memory>
#ifdef _OPENMP
#include <omp.h>
#endif

#define REF_API extern "C" __attribute__((visibility("default")))

namespace {
struct RefLaser {
    ros::shim::Context ctx;
    std::unique_ptr<aslam::LandMarks> node;
    RefLaser() {
        ros::shim::Use use(&ctx);
        node.reset(new aslam::LandMarks());
        // sensor_landmark.cpp:87,98 call `bearing2pose(theta, msg->ranges[theta++])`. The order in which the two
        // arguments are evaluated is unspecified; GCC (the compiler of every ROS build, and of this one) evaluates
        // them right to left, so ray t is projected with the table entry t + 1, and ray 359 reads cos_map[360] /
        // sin_map[360] -- one past the 360 entries initialize() pushed (:50-57), inside the vector's spare capacity.
        // A fresh heap holds 0.0f there; make that explicit so that _ref is deterministic.
        node->cos_map.reserve(361);
        node->sin_map.reserve(361);
        node->cos_map.data()[360] = 0.0f;
        node->sin_map.data()[360] = 0.0f;
    }
    // One LaserScan through LandMarks::callback. Returns the number of published landmarks, or -1 without calling
    // the node when its `assert(theta < 359)` (sensor_landmark.cpp:82) would abort the process: first and last ray
    // within MIN_DIST_THRESH make the first-cluster loop (:73-79) run to theta = 360 (p1 == p2 after each pass).
    int scan(const float* ranges, float* out_r, float* out_b, int cap) {
        Point p1 = node->bearing2pose(0, ranges[0]);
        Point p2 = node->bearing2pose(359, ranges[359]);
        if (p1.distance(p2) < MIN_DIST_THRESH) return -1;
        auto msg = std::make_shared<sensor_msgs::LaserScan>();
        msg->ranges.assign(ranges, ranges + 360);
        ros::shim::Use use(&ctx);
        ctx.dispatch<sensor_msgs::LaserScan>("/laser/scan", msg);
        const auto* m = ctx.last_published<awesome_slam_msgs::Landmarks>("out/landmarks/sensor");
        const int n = m ? (int)m->x.size() : 0;
        for (int i = 0; i < n && i < cap; ++i) {
            out_r[i] = (float)m->x[i];  // the message doubles hold float values (LaserData members)
            out_b[i] = (float)m->y[i];
        }
        return n;
    }
};
}  // namespace

REF_API void* ref_laser_create() { return new RefLaser; }
REF_API void ref_laser_destroy(void* p) { delete (RefLaser*)p; }
REF_API int ref_laser_scan(void* p, const float* ranges, float* out_r, float* out_b, int cap) {
    return ((RefLaser*)p)->scan(ranges, out_r, out_b, cap);
}
// nscan scans [nscan][360] -> out_r / out_b [nscan][cap], counts [nscan]; one node per thread. Returns seconds.
REF_API double ref_laser_extract(const float* ranges, int nscan, int cap, float* out_r, float* out_b, int* counts,
                                 int nthreads) {
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel
    {
        RefLaser node;
#pragma omp for schedule(dynamic, 16)
        for (int s = 0; s < nscan; ++s)
            counts[s] = node.scan(ranges + (size_t)s * 360, out_r + (size_t)s * cap, out_b + (size_t)s * cap, cap);
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}
