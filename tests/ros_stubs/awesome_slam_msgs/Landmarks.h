#pragma once
#include <memory>
#include <vector>
namespace awesome_slam_msgs {
struct Landmarks {  // awesome_slam_msgs/msg/Landmarks.msg:1-2
    std::vector<double> x, y;
    typedef std::shared_ptr<const Landmarks> ConstPtr;
};
}  // namespace awesome_slam_msgs
