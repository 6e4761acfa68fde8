// TEST INFRASTRUCTURE ONLY -- builds the reference's OWN UKF node (src/ukf/ukf.cpp with ukf.h, common.h, tools.h,
// structures.h, config.h, compiled from where they lie under /root/reference, unedited) into
// oracle/_ref/libref_ukf.so, against the Eigen / ROS stand-ins of oracle/ref_shim/include. See README.md here.
#define main aslam_ref_ukf_node_main  // ukf.cpp:394 defines main(); it is never called
#include <ukf/ukf.cpp>
#undef main

#define REF_NODE aslam::UKFSlam
#define REF_HAS_A 0
#include "filter_driver.inc"
