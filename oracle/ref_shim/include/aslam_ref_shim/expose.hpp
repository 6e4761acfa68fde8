// TEST INFRASTRUCTURE ONLY. The reference's node classes keep their state and their callbacks private
// (ekf.h:77-131, ukf.h:92-143, sensor_landmark.h:56-84) and the reference has no accessor for X, P, Z, N or the
// wait-list. The drivers in oracle/ref_shim/ need to READ those members to compare them bit for bit, so the build
// of oracle/_ref defines ASLAM_REF_EXPOSE and this header -- included at the end of every Eigen stand-in, i.e.
// after the last standard header has been seen (the reference's .cpp files include nothing after their own .h) --
// turns the keyword off for the class definitions that follow. The reference text itself is not edited.
#ifdef ASLAM_REF_EXPOSE
#ifndef ASLAM_REF_EXPOSE_DONE
#define ASLAM_REF_EXPOSE_DONE
#define private public
#endif
#endif
