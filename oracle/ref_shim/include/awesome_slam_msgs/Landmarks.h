// TEST INFRASTRUCTURE ONLY -- awesome_slam_msgs/msg/Landmarks.msg:1-2 (float64[] x, float64[] y) as ROS would
// generate it.
#pragma once
#include <memory>
#include <vector>
namespace awesome_slam_msgs {
struct Landmarks {
    std::vector<double> x, y;
    typedef std::shared_ptr<const Landmarks> ConstPtr;
};
}  // namespace awesome_slam_msgs
