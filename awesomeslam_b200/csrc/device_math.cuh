// Scalar device functions that reproduce the reference's FP32 truncation sites exactly
// (SURVEY.md section 8a-casts). The whole library is compiled with -fmad=false, so `a*b+c`
// never contracts; hot loops ask for FMAs explicitly with fma(). The intrinsics below make
// the rounding points explicit where a float result is later compared or re-widened.
// Reference paths are relative to awesome_slam/ in the reference tree.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/aslam_c.h"

namespace aslam {

// 1 / x for a well-scaled positive x (the determinant of an innovation covariance): the hardware seed
// (MUFU.RCP64H, about 20 bits) and two Newton steps -- full double precision to the last bit or two, without the
// special-case path of the IEEE division. Only used where the result feeds FP64 arithmetic under the 1e-9 contract,
// never at one of the reference's float cast sites.
__device__ __forceinline__ double fast_rcp(double x) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    r = fma(fma(-x, r, 1.0), r, r);
    r = fma(fma(-x, r, 1.0), r, r);
    return r;
}

// tools.h:44-50 normalizeAngle: float in, float out, fmodf (exact on the device), 2*PI is a
// float product.
__device__ __forceinline__ float normalize_angle(float theta, float pi) {
    const float two_pi = __fmul_rn(2.0f, pi);
    // fmodf is exact, so its two common cases are written out: |theta| < 2pi returns theta itself and
    // 2pi <= |theta| < 4pi returns |theta| - 2pi, a float subtraction that is exact (Sterbenz), with
    // theta's sign. Everything else (and NaN/Inf) takes the library routine.
    const float a = fabsf(theta);
    float ret;
    if (a < two_pi)
        ret = theta;
    else if (a < __fmul_rn(2.0f, two_pi))
        ret = copysignf(__fsub_rn(a, two_pi), theta);
    else
        ret = fmodf(theta, two_pi);
    ret = ret > pi ? __fsub_rn(ret, two_pi) : ret;
    ret = ret < -pi ? __fadd_rn(ret, two_pi) : ret;
    return ret;
}

// tools.h:53-59 eulerDistance on two Points whose coordinates are float-exact doubles
// (structures.h:52): dx, dy rounded to float; sqrtf(dx*dx + dy*dy) with separate roundings.
// The reference subtracts the float-exact doubles and rounds the difference to float; that is the
// float subtraction: the double difference is exact whenever the exponents are within 29 bits, and
// beyond that both roundings return the larger operand.
__device__ __forceinline__ float euler_distance(float ax, float ay, float bx, float by) {
    const float dx = __fsub_rn(ax, bx);
    const float dy = __fsub_rn(ay, by);
    return __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
}

// structures.h:104-111 LaserData::toPoint(Z): double evaluation, float store.
__device__ __forceinline__ float2 to_point(float range, float bearing, double z0, double z1, double z2) {
    double s, c;
    sincos(__dadd_rn(z2, (double)bearing), &s, &c);
    float2 p;
    p.x = (float)__dadd_rn(z0, __dmul_rn((double)range, c));
    p.y = (float)__dadd_rn(z1, __dmul_rn((double)range, s));
    return p;
}

// common.h:46-65 stateTransitionFunction, non-augmented part, rows 0..2 (the only rows that
// change). vx, az, dt are the float arguments of slam() (ekf.cpp:293, ukf.cpp:260).
__device__ __forceinline__ void motion_model(double& x, double& y, double& th, float vx, float az, float dt) {
    const double th0 = th;
    const float azdt = __fmul_rn(az, dt);  // common.h:55-56,64: f32 product, then widened
    if (fabs((double)az) > 0.001) {        // common.h:52
        const float r = __fdiv_rn(vx, az);  // common.h:54
        double s0, c0, s1, c1;
        sincos(th0, &s0, &c0);
        sincos(__dadd_rn(th0, (double)azdt), &s1, &c1);
        x = __dadd_rn(x, __dmul_rn((double)r, __dadd_rn(-s0, s1)));
        y = __dadd_rn(y, __dmul_rn((double)r, __dsub_rn(c0, c1)));
    } else {
        const float vxdt = __fmul_rn(vx, dt);  // common.h:60-61
        double s0, c0;
        sincos(th0, &s0, &c0);
        x = __dadd_rn(x, __dmul_rn((double)vxdt, c0));
        y = __dadd_rn(y, __dmul_rn((double)vxdt, s0));
    }
    th = __dadd_rn(th0, (double)azdt);
}

// common.h:46-75 for one (augmented) sigma point: rows 0..2 of the column; `na` = point(N).
__device__ __forceinline__ void motion_model_aug(double& x, double& y, double& th, double na, float vx, float az,
                                                 float dt) {
    const double th0 = th;
    double s0, c0;
    sincos(th0, &s0, &c0);
    const float azdt = __fmul_rn(az, dt);
    if (fabs((double)az) > 0.001) {
        const float r = __fdiv_rn(vx, az);
        double s1, c1;
        sincos(__dadd_rn(th0, (double)azdt), &s1, &c1);
        x = __dadd_rn(x, __dmul_rn((double)r, __dadd_rn(-s0, s1)));
        y = __dadd_rn(y, __dmul_rn((double)r, __dsub_rn(c0, c1)));
    } else {
        const float vxdt = __fmul_rn(vx, dt);
        x = __dadd_rn(x, __dmul_rn((double)vxdt, c0));
        y = __dadd_rn(y, __dmul_rn((double)vxdt, s0));
    }
    th = __dadd_rn(th0, (double)azdt);
    // common.h:69-71: 0.5 * dt * dt * point(N) * cos(point(2)), all double, left to right
    const double h = __dmul_rn(__dmul_rn(0.5, (double)dt), (double)dt);
    x = __dadd_rn(x, __dmul_rn(__dmul_rn(h, na), c0));
    y = __dadd_rn(y, __dmul_rn(__dmul_rn(h, na), s0));
    th = __dadd_rn(th, __dmul_rn(h, (double)az));  // `az`, not point(N+1), as written
}

// ekf.cpp:206-211: the two information-bearing entries of A. vx, az are the message DOUBLES.
__device__ __forceinline__ void motion_jacobian(double vx, double az, float dt, double yaw, double& a00,
                                                double& a10) {
    if (vx != 0.0 && az != 0.0) {
        const float delta_theta = (float)__dmul_rn(az, (double)dt);
        const float r = (float)__ddiv_rn(vx, az);
        double s0, c0, s1, c1;
        sincos(yaw, &s0, &c0);
        sincos(__dadd_rn(yaw, (double)delta_theta), &s1, &c1);
        a00 = __dmul_rn((double)r, __dadd_rn(-c0, c1));
        a10 = __dmul_rn((double)r, __dadd_rn(-s0, s1));
    }
}

// ekf.cpp:121-132 updateH for one landmark (lx, ly) seen from (x, y): hyp and dist are floats.
//   row range   (3+i): [ a,  c,  0, ... -a, -c ]
//   row bearing (4+i): [ b,  d, -1, ... -b, -d ]
struct HEntries {
    double a, b, c, d;
};
__device__ __forceinline__ HEntries jacobian_entries(double x, double y, double lx, double ly) {
    const double dx = __dsub_rn(lx, x), dy = __dsub_rn(ly, y);
    const float hyp = (float)__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
    const float dist = __fsqrt_rn(hyp);
    const double nx = __dadd_rn(-lx, x), ny = __dadd_rn(-ly, y);
    HEntries h;
    h.a = __ddiv_rn(nx, (double)dist);   // H(3+i,0); H(3+i,3+i) = -a
    h.b = __ddiv_rn(-ny, (double)hyp);   // H(4+i,0); H(4+i,3+i) = -b
    h.c = __ddiv_rn(ny, (double)dist);   // H(3+i,1); H(3+i,4+i) = -c
    h.d = __ddiv_rn(nx, (double)hyp);    // H(4+i,1); H(4+i,4+i) = -d
    return h;
}

// common.h:86-87 measurementFunction for one landmark (bearing left un-wrapped, as written).
__device__ __forceinline__ void measure_landmark(double x, double y, double th, double lx, double ly,
                                                 double& range, double& bearing) {
    const double dx = __dsub_rn(lx, x), dy = __dsub_rn(ly, y);
    range = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    bearing = __dsub_rn(atan2(dy, dx), th);
}

}  // namespace aslam
