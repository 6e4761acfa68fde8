#pragma once
#include <memory>
namespace nav_msgs {
struct Odometry {
    struct { struct { struct { double x, y, z; } position; struct { double x, y, z, w; } orientation; } pose; } pose;
    struct { struct { struct { double x, y, z; } linear, angular; } twist; } twist;
    typedef std::shared_ptr<const Odometry> ConstPtr;
};
}  // namespace nav_msgs
