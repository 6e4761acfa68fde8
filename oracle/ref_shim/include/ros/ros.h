// TEST INFRASTRUCTURE ONLY -- stand-in for <ros/ros.h>, see ../aslam_ref_shim/ros_context.hpp.
#pragma once
#include "../aslam_ref_shim/ros_context.hpp"
