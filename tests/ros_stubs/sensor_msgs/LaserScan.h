#pragma once
#include <memory>
#include <vector>
namespace sensor_msgs {
struct LaserScan {
    std::vector<float> ranges;
    typedef std::shared_ptr<const LaserScan> ConstPtr;
};
}  // namespace sensor_msgs
