// TEST INFRASTRUCTURE ONLY -- included by ekf.h:44 / ukf.h:44 but never used by the reference.
#pragma once
namespace geometry_msgs {
struct Twist {
    struct Vec3 { double x = 0, y = 0, z = 0; } linear, angular;
};
}  // namespace geometry_msgs
