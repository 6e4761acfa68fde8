// TEST INFRASTRUCTURE ONLY -- builds the reference's OWN EKF node (src/ekf/ekf.cpp with ekf.h, common.h, tools.h,
// structures.h, config.h, compiled from where they lie under /root/reference, unedited) into
// oracle/_ref/libref_ekf.so, against the Eigen / ROS stand-ins of oracle/ref_shim/include. See README.md here.
#define main aslam_ref_ekf_node_main  // ekf.cpp:313 defines main(); it is never called
#include <ekf/ekf.cpp>
#undef main

#define REF_NODE aslam::EKFSlam
#define REF_HAS_A 1
#include "filter_driver.inc"
