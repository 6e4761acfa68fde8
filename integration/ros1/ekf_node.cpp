// ROS1 (Noetic) adapter for the filter nodes. NOT built in this repository's image (no ROS here);
// it is the file a maintainer drops next to awesome_slam/src/ekf/ekf.cpp so that `rosrun awesome_slam ekf` (or `ukf`,
// -DASLAM_NODE_UKF) runs on libaslam_b200.so with the reference's topics, message types and call order unchanged:
//   sub  /odom                  nav_msgs/Odometry            (src/ekf/ekf.cpp:41)   -> aslam::EKFSlam::cbOdom
//   sub  out/landmarks/sensor   awesome_slam_msgs/Landmarks  (:42)                  -> cbSensorLandmark
//   pub  out/landmarks/kalman   awesome_slam_msgs/Landmarks  (:43, 97-98)
// The filter mathematics is include/aslam_filters.hpp over the C ABI (include/aslam_c.h); nothing here computes.
#include <awesome_slam_msgs/Landmarks.h>
#include <nav_msgs/Odometry.h>
#include <ros/ros.h>

#include <algorithm>

#include "aslam_filters.hpp"

#ifdef ASLAM_NODE_UKF
using Filter = aslam::UKFSlam;
static const char* kNodeName = "aslam_ukf";
#else
using Filter = aslam::EKFSlam;
static const char* kNodeName = "aslam_ekf";
#endif

class FilterNode {
  public:
    FilterNode() : nh_(ros::NodeHandle()) {
        sub_odom_ = nh_.subscribe("/odom", 1, &FilterNode::cbOdom, this);
        sub_sensor_ = nh_.subscribe("out/landmarks/sensor", 1, &FilterNode::cbSensorLandmark, this);
        pub_ = nh_.advertise<awesome_slam_msgs::Landmarks>("out/landmarks/kalman", 1);
        last_time_ = static_cast<float>(ros::Time::now().toSec());  // ekf.cpp:69
    }

  private:
    void cbSensorLandmark(const awesome_slam_msgs::Landmarks::ConstPtr& msg) {  // ekf.cpp:102-114
        aslam::Landmarks lm;
        lm.x.assign(msg->x.begin(), msg->x.end());
        lm.y.assign(msg->y.begin(), msg->y.end());
        filter_.cbSensorLandmark(lm);
    }
    void cbOdom(const nav_msgs::Odometry::ConstPtr& msg) {  // ekf.cpp:74-99
        const double now = ros::Time::now().toSec();
        const float delta_time = static_cast<float>(std::min(now - last_time_, 1.0));  // ekf.cpp:80
        last_time_ = static_cast<float>(now);                                           // ekf.cpp:81
        aslam::Odometry od;
        od.position.x = msg->pose.pose.position.x;
        od.position.y = msg->pose.pose.position.y;
        od.orientation.w = msg->pose.pose.orientation.w;
        od.orientation.x = msg->pose.pose.orientation.x;
        od.orientation.y = msg->pose.pose.orientation.y;
        od.orientation.z = msg->pose.pose.orientation.z;
        od.linear_x = msg->twist.twist.linear.x;
        od.angular_z = msg->twist.twist.angular.z;
        aslam::Landmarks out;
        if (!filter_.cbOdom(od, delta_time, &out)) return;  // gated until the first sensor message (ekf.cpp:76)
        awesome_slam_msgs::Landmarks lm;
        lm.x = out.x;
        lm.y = out.y;
        pub_.publish(lm);
    }

    ros::NodeHandle nh_;
    ros::Subscriber sub_odom_, sub_sensor_;
    ros::Publisher pub_;
    Filter filter_;
    float last_time_ = 0.0f;
};

int main(int argc, char** argv) {  // ekf.cpp:313-326
    ros::init(argc, argv, kNodeName);
    ros::Time::init();
    ros::Rate rate(1);  // FREQ (config.h:42)
    FilterNode node;
    while (ros::ok()) {
        ros::spinOnce();
        rate.sleep();
    }
    return 0;
}
