// TEST INFRASTRUCTURE ONLY -- the one field of sensor_msgs/LaserScan the reference reads (sensor_landmark.cpp:62-98).
#pragma once
#include <memory>
#include <vector>
namespace sensor_msgs {
struct LaserScan {
    std::vector<float> ranges;
    typedef std::shared_ptr<const LaserScan> ConstPtr;
};
}  // namespace sensor_msgs
