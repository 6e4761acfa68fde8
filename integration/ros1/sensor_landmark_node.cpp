// ROS1 (Noetic) adapter for the laser front end. NOT built in this repository's image (no ROS).
//   sub  /laser/scan            sensor_msgs/LaserScan        (src/sensor_landmark/sensor_landmark.cpp:40)
//   pub  out/landmarks/sensor   awesome_slam_msgs/Landmarks  (:41, 127-132)
// The clustering, circle test and hyper-fit run in libaslam_b200.so (aslam_laser_extract, include/aslam_c.h).
#include <awesome_slam_msgs/Landmarks.h>
#include <ros/ros.h>
#include <sensor_msgs/LaserScan.h>

#include "aslam_c.h"

class LandMarksNode {
  public:
    LandMarksNode() : nh_(ros::NodeHandle()) {
        if (aslam_laser_create(nullptr, 1, 0, &gpu_) != ASLAM_OK) throw std::runtime_error(aslam_last_error());
        sub_ = nh_.subscribe("/laser/scan", 10, &LandMarksNode::callback, this);
        pub_ = nh_.advertise<awesome_slam_msgs::Landmarks>("out/landmarks/sensor", 10);
    }
    ~LandMarksNode() { aslam_laser_destroy(gpu_); }

  private:
    void callback(const sensor_msgs::LaserScan::ConstPtr& msg) {  // sensor_landmark.cpp:59-133
        if (msg->ranges.size() < ASLAM_LASER_RAYS) return;
        float range[ASLAM_LASER_MAX_LANDMARKS], bearing[ASLAM_LASER_MAX_LANDMARKS];
        int32_t count = 0;
        if (aslam_laser_extract(gpu_, msg->ranges.data(), 1, range, bearing, &count) != ASLAM_OK) return;
        if (count < 0) return;  // the reference's assert(theta < 359) case (:82)
        awesome_slam_msgs::Landmarks lm;
        lm.x.assign(range, range + count);
        lm.y.assign(bearing, bearing + count);
        pub_.publish(lm);
    }
    ros::NodeHandle nh_;
    ros::Subscriber sub_;
    ros::Publisher pub_;
    aslam_laser* gpu_ = nullptr;
};

int main(int argc, char** argv) {  // sensor_landmark.cpp:300-313
    ros::init(argc, argv, "aslam_sensor_landmark");
    ros::Time::init();
    ros::Rate rate(1);
    LandMarksNode node;
    while (ros::ok()) {
        ros::spinOnce();
        rate.sleep();
    }
    return 0;
}
