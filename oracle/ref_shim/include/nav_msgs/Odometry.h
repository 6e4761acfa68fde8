// TEST INFRASTRUCTURE ONLY -- the fields of nav_msgs/Odometry the reference reads (ekf.cpp:94,139-142,206-209).
#pragma once
#include <memory>
namespace nav_msgs {
struct Odometry {
    struct Vec3 { double x = 0, y = 0, z = 0; };
    struct Quat { double x = 0, y = 0, z = 0, w = 1; };
    struct Pose { Vec3 position; Quat orientation; };
    struct PoseWithCovariance { Pose pose; };
    struct Twist { Vec3 linear, angular; };
    struct TwistWithCovariance { Twist twist; };
    PoseWithCovariance pose;
    TwistWithCovariance twist;
    typedef std::shared_ptr<const Odometry> ConstPtr;
};
}  // namespace nav_msgs
