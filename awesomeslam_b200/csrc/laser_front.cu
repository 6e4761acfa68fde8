// Laser front end on the GPU: the reference's sensor_landmark node
// (awesome_slam/src/sensor_landmark/sensor_landmark.cpp:59-298) for a batch of LaserScans -- bearing2pose, gap
// clustering at MIN_DIST_THRESH, the inscribed-angle circle test (Xavier et al. 2005, :147-186), the Chernov
// hyper-fit (:192-298) and Circle::toLaserData (include/awesome_slam/structures.h:126-133) -- producing the
// (range, bearing) lists the filters consume (cbSensorLandmark, src/ekf/ekf.cpp:102-114).
//
// One warp per scan. Lanes project the 360 rays, mark the cluster breaks and pair consecutive breaks into candidate
// clusters with ballots (the reference's loop drops the point that opens every cluster after the first and never
// evaluates the trailing run: reproduced as written); then every candidate cluster is classified and fitted by one lane, and the accepted
// ones are compacted in scan order with a ballot. Float sites of the reference stay float (explicit _rn intrinsics;
// the library is built with -fmad=false); the 4 x 4 eigen-problems are cyclic Jacobi in registers.
// The sin/cos tables are built on the HOST exactly as LandMarks::initialize does (:50-57) and uploaded.
#include <cuda_runtime.h>

#include <cmath>
#include <new>

#include "device_math.cuh"
#include "internal.hpp"
#include "nvtx_range.hpp"

namespace aslam {

namespace {

constexpr int RAYS = 360;
constexpr int MAXC = 96;   // candidate clusters per scan (a cluster needs > 3 points: at most 360 / 5)
constexpr int LWPB = 4;    // warps (scans) per CTA

struct LaserParams {
    float thresh, std_max, min_mean, max_mean;
    int min_pts;
    float c0, s0, c359, s359;  // table entries 0 and 359 as initialize() builds them: the wrap test (:63-64)
};

struct __align__(16) ScanSlot {
    float px[RAYS], py[RAYS];
    float ang[RAYS];  // inscribed angles of the cluster a ray belongs to (clusters are disjoint ray ranges)
    short cstart[MAXC], clen[MAXC];
    int ncand;
};

// The points of one cluster, as the classification and the fit read them: from the scan's slot in shared memory, or
// recomputed from the ranges in global memory (the same float product as bearing2pose, :136-142).
struct SlotPoints {
    const float *px, *py;
    __device__ __forceinline__ float x(int i) const { return px[i]; }
    __device__ __forceinline__ float y(int i) const { return py[i]; }
};
struct RangePoints {
    const float *r, *c, *s;
    __device__ __forceinline__ float x(int i) const { return __fmul_rn(__ldg(r + i), __ldg(c + i)); }
    __device__ __forceinline__ float y(int i) const { return __fmul_rn(__ldg(r + i), __ldg(s + i)); }
};

// cyclic Jacobi, symmetric 4 x 4: A = V diag(w) V'
__device__ __forceinline__ void jacobi4(double (&A)[4][4], double (&V)[4][4]) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) V[i][j] = i == j ? 1.0 : 0.0;
    // classical cyclic Jacobi: the rotated entry is set to exactly 0; from the 4th sweep on an entry too small to
    // change either of its diagonal neighbours is dropped; ends when the off-diagonal part is exactly zero
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0;
#pragma unroll
        for (int i = 0; i < 3; ++i)
#pragma unroll
            for (int j = i + 1; j < 4; ++j) off += fabs(A[i][j]);
        if (off == 0.0) break;
#pragma unroll
        for (int p = 0; p < 3; ++p)
#pragma unroll
            for (int q = p + 1; q < 4; ++q) {
                if (A[p][q] != 0.0) {
                    const double g = 100.0 * fabs(A[p][q]);
                    if (sweep > 3 && fabs(A[p][p]) + g == fabs(A[p][p]) && fabs(A[q][q]) + g == fabs(A[q][q])) {
                        A[p][q] = 0.0;
                        A[q][p] = 0.0;
                    } else {
                        const double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
                        const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                        const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const double akp = A[k][p], akq = A[k][q];
                            A[k][p] = c * akp - s * akq;
                            A[k][q] = s * akp + c * akq;
                        }
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const double apk = A[p][k], aqk = A[q][k];
                            A[p][k] = c * apk - s * aqk;
                            A[q][k] = s * apk + c * aqk;
                        }
                        A[p][q] = 0.0;
                        A[q][p] = 0.0;
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const double vkp = V[k][p], vkq = V[k][q];
                            V[k][p] = c * vkp - s * vkq;
                            V[k][q] = s * vkp + c * vkq;
                        }
                    }
                }
            }
    }
}

// sensor_landmark.cpp:147-186 on the n points of one cluster
template <typename PTS>
__device__ __forceinline__ bool classify(const PTS& pts, const LaserParams& lp, int n, float* ang) {
    // ang: n floats of scratch owned by this cluster (the reference keeps the angles in a vector, :152-171)
    const float x1 = pts.x(0), y1 = pts.y(0), x2 = pts.x(n - 1), y2 = pts.y(n - 1);
    const float c = euler_distance(x1, y1, x2, y2);
    const float cc = __fmul_rn(c, c);
    float sum = 0.0f;
    for (int i = 1; i < n - 1; ++i) {
        const float x = pts.x(i), y = pts.y(i);
        const float a = euler_distance(x1, y1, x, y), b = euler_distance(x2, y2, x, y);
        const float num = __fsub_rn(__fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b)), cc);
        const float den = __fmul_rn(__fmul_rn(2.0f, a), b);
        // acosf of the reference's libm, taken as the correctly rounded value of the double routine
        const float angle = (float)acos((double)__fdiv_rn(num, den));
        ang[i] = angle;
        sum = __fadd_rn(sum, angle);
    }
    const float mean = __fdiv_rn(sum, (float)(n - 2));
    float variance = 0.0f;
    for (int i = 1; i < n - 1; ++i) {
        const double d = (double)__fsub_rn(ang[i], mean);
        variance = (float)__dadd_rn((double)variance, __dmul_rn(d, d));  // pow(float, 2) is a double (:174)
    }
    const float sd = (float)sqrt(__ddiv_rn((double)variance, 5.0));  // (:177)
    return sd < lp.std_max && mean > lp.min_mean && mean < lp.max_mean;
}

// sensor_landmark.cpp:192-298 (hyper-fit) + structures.h:126-133 on the n points of one cluster
template <typename PTS>
__device__ __forceinline__ void fit(const PTS& pts, int n, float& out_range, float& out_bearing) {
    float x_mean = 0.0f, y_mean = 0.0f, z_mean = 0.0f;
    for (int i = 0; i < n; ++i) {
        x_mean = __fadd_rn(x_mean, pts.x(i));
        y_mean = __fadd_rn(y_mean, pts.y(i));
    }
    x_mean = __fdiv_rn(x_mean, (float)n);
    y_mean = __fdiv_rn(y_mean, (float)n);
    for (int i = 0; i < n; ++i) {
        const float x = __fsub_rn(pts.x(i), x_mean), y = __fsub_rn(pts.y(i), y_mean);
        z_mean = __fadd_rn(z_mean, __fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)));
    }
    z_mean = __fdiv_rn(z_mean, (float)n);
    // G = Z'Z for Z = [z x y 1]
    double G[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) G[a][b] = 0.0;
    for (int i = 0; i < n; ++i) {
        const float x = __fsub_rn(pts.x(i), x_mean), y = __fsub_rn(pts.y(i), y_mean);
        const float z = __fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y));
        const double row[4] = {(double)z, (double)x, (double)y, 1.0};
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) G[a][b] = __dadd_rn(G[a][b], __dmul_rn(row[a], row[b]));
    }
    double V[4][4];
    jacobi4(G, V);
    double sv[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) sv[i] = sqrt(G[i][i] > 0.0 ? G[i][i] : 0.0);
    double sv_min = sv[0];
#pragma unroll
    for (int i = 1; i < 4; ++i) sv_min = sv[i] < sv_min ? sv[i] : sv_min;
    double A[4];
    if (sv_min > 10e-12) {  // (:256)
        const double hz = (double)__fmul_rn(-2.0f, z_mean);  // H^-1(3,3) (:245)
        // Y = V Sigma V'
        double Y[4][4], T[4][4], Q[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                double acc = 0.0;
#pragma unroll
                for (int k = 0; k < 4; ++k) acc += V[a][k] * sv[k] * V[b][k];
                Y[a][b] = acc;
            }
        // T = Y H^-1 with H^-1 = [0 0 0 1/2; 0 1 0 0; 0 0 1 0; 1/2 0 0 hz]
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            T[a][0] = Y[a][3] * 0.5;
            T[a][1] = Y[a][1];
            T[a][2] = Y[a][2];
            T[a][3] = Y[a][0] * 0.5 + Y[a][3] * hz;
        }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b <= a; ++b) {  // the eigen-solver reads the lower triangle
                double acc = 0.0;
#pragma unroll
                for (int k = 0; k < 4; ++k) acc += T[a][k] * Y[k][b];
                Q[a][b] = acc;
                Q[b][a] = acc;
            }
        double E[4][4];
        jacobi4(Q, E);
        // eigenvector of the smallest positive eigenvalue (< 99999), first one in ascending order (:259-283)
        int col = -1;
        double best = 99999.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const double ev = Q[i][i];
            if (ev > 0.0 && ev < best) {
                best = ev;
                col = i;
            }
        }
        if (col < 0) {  // no positive eigenvalue: the reference keeps smallest_id = 0, the most negative one
            double lo = Q[0][0];
            col = 0;
#pragma unroll
            for (int i = 1; i < 4; ++i)
                if (Q[i][i] < lo) {
                    lo = Q[i][i];
                    col = i;
                }
        }
        // A = Y^-1 A_star = V Sigma^-1 V' A_star
        double t4[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            double d = 0.0;
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                double e = E[a][0];
                if (col == 1) e = E[a][1];
                if (col == 2) e = E[a][2];
                if (col == 3) e = E[a][3];
                d += V[a][k] * e;
            }
            t4[k] = d / sv[k];
        }
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            double acc = 0.0;
#pragma unroll
            for (int k = 0; k < 4; ++k) acc += V[a][k] * t4[k];
            A[a] = acc;
        }
    } else {  // A = V.col(3): the singular vector of the smallest singular value
        int im = 0;
#pragma unroll
        for (int i = 1; i < 4; ++i)
            if (sv[i] < sv[im]) im = i;
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            double e = V[a][0];
            if (im == 1) e = V[a][1];
            if (im == 2) e = V[a][2];
            if (im == 3) e = V[a][3];
            A[a] = e;
        }
    }
    const float a = (float)(-A[1] / (2.0 * A[0]));
    const float b = (float)(-A[2] / (2.0 * A[0]));
    const double cx = (double)__fadd_rn(a, x_mean), cy = (double)__fadd_rn(b, y_mean);
    out_range = (float)sqrt(__dadd_rn(__dmul_rn(cx, cx), __dmul_rn(cy, cy)));
    out_bearing = (float)atan2(cy, cx);
}

__global__ void __launch_bounds__(LWPB * 32) laser_extract_kernel(const float* __restrict__ ranges, uint32_t nscan,
                                                                 const float* __restrict__ cos_map,
                                                                 const float* __restrict__ sin_map, LaserParams lp,
                                                                 uint2* __restrict__ work,
                                                                 uint32_t* __restrict__ work_count,
                                                                 int32_t* __restrict__ counts) {
    __shared__ ScanSlot slots[LWPB];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t scan = blockIdx.x * LWPB + warp;
    if (scan >= nscan) return;  // whole warp; no CTA barrier in this kernel
    ScanSlot& s = slots[warp];
    const float* r = ranges + (size_t)scan * RAYS;
    // bearing2pose (:136-142). cos_map / sin_map are the POSE tables: entry t is what the reference multiplies ray t
    // with inside its clustering loop, i.e. initialize()'s entry t + 1 (see aslam_laser_create).
    for (int t = lane; t < RAYS; t += 32) {
        const float d = r[t];
        s.px[t] = __fmul_rn(d, cos_map[t]);
        s.py[t] = __fmul_rn(d, sin_map[t]);
    }
    __syncwarp();
    // first and last ray on the same object: the reference's assert(theta < 359) fires (:82). This test (:63-64)
    // projects rays 0 and 359 with their own table entries.
    if (euler_distance(__fmul_rn(r[0], lp.c0), __fmul_rn(r[0], lp.s0), __fmul_rn(r[RAYS - 1], lp.c359),
                       __fmul_rn(r[RAYS - 1], lp.s359)) < lp.thresh) {
        if (lane == 0) counts[scan] = -1;
        return;
    }
    // (:85-124) A break at ray t (the step from t-1 to t is not below the threshold; NaN counts) closes the cluster
    // [start, t) and the next one starts at t + 1: the candidates are the pairs of consecutive breaks (b, t) with
    // t - b - 1 > MIN_CLUSTER_POINTS (first cluster: start = 0). Every lane looks at one ray per round; the previous
    // break comes from the ballot of its round or from the running last break of the earlier rounds; candidates are
    // appended in scan order (ballot + prefix count).
    {
        int last_break = -1, nc = 0;  // warp-uniform
        for (int base = 1; base < RAYS; base += 32) {
            const int t = base + lane;
            const bool brk = t < RAYS && !(euler_distance(s.px[t - 1], s.py[t - 1], s.px[t], s.py[t]) < lp.thresh);
            const unsigned m = __ballot_sync(0xffffffffu, brk);
            const unsigned below = m & ((1u << lane) - 1u);
            const int prev = below ? base + (31 - __clz(below)) : last_break;
            const int start = prev + 1;  // prev = -1: the first cluster starts at ray 0
            const bool cand = brk && (t - start) > lp.min_pts;
            const unsigned cm = __ballot_sync(0xffffffffu, cand);
            const int slot = nc + __popc(cm & ((1u << lane) - 1u));
            if (cand && slot < MAXC) {
                s.cstart[slot] = (short)start;
                s.clen[slot] = (short)(t - start);
            }
            nc += __popc(cm);
            if (m) last_break = base + (31 - __clz(m));
        }
        if (lane == 0) s.ncand = nc < MAXC ? nc : MAXC;
    }
    __syncwarp();
    // Classification: one lane per candidate. The accepted clusters keep their scan order (ballot + prefix count) and
    // go, with the output slot they will fill, to a batch-wide dense work list; the fit runs in laser_fit_kernel.
    const int nc = s.ncand;
    int total = 0;
    for (int base = 0; base < nc; base += 32) {
        const int cnd = base + lane;
        bool ok = false;
        int s0 = 0, n = 0;
        if (cnd < nc) {
            s0 = s.cstart[cnd];
            n = s.clen[cnd];
            ok = classify(SlotPoints{s.px + s0, s.py + s0}, lp, n, s.ang + s0);
        }
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        const int slot = total + __popc(m & ((1u << lane) - 1u));
        const bool emit = ok && slot < (int)ASLAM_LASER_MAX_LANDMARKS;
        const unsigned em = __ballot_sync(0xffffffffu, emit);
        if (em) {
            uint32_t wbase = 0;
            if (lane == 0) wbase = atomicAdd(work_count, (uint32_t)__popc(em));
            wbase = __shfl_sync(0xffffffffu, wbase, 0);
            if (emit)
                work[wbase + __popc(em & ((1u << lane) - 1u))] =
                    make_uint2(scan * ASLAM_LASER_MAX_LANDMARKS + (uint32_t)slot, (uint32_t)s0 | ((uint32_t)n << 16));
        }
        total += __popc(m);
    }
    if (lane == 0) counts[scan] = total;
}

// Circle fit of the accepted clusters, one thread per cluster of the whole batch (every lane busy; the order of the
// work list does not matter: each entry names the output slot it fills).
__global__ void __launch_bounds__(128) laser_fit_kernel(const float* __restrict__ ranges, const float* __restrict__ cos_map,
                                                        const float* __restrict__ sin_map,
                                                        const uint2* __restrict__ work,
                                                        const uint32_t* __restrict__ work_count,
                                                        float* __restrict__ out_range, float* __restrict__ out_bearing) {
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= *work_count) return;
    const uint2 w = work[idx];
    const uint32_t scan = w.x / ASLAM_LASER_MAX_LANDMARKS;
    const int s0 = (int)(w.y & 0xffffu), n = (int)(w.y >> 16);
    float rg, bg;
    fit(RangePoints{ranges + (size_t)scan * RAYS + s0, cos_map + s0, sin_map + s0}, n, rg, bg);
    out_range[w.x] = rg;
    out_bearing[w.x] = bg;
}

// the Gazebo LDS model as arithmetic -- one thread per ray: ray k of a scan leaves the sensor at
// yaw + k degrees, hits the nearest cylinder (radius r) of the map, returns inside [rmin, rmax] get sigma * noise
// added, everything else is +inf (what Gazebo publishes for "no return").
__global__ void __launch_bounds__(256) laser_synth_kernel(const double* __restrict__ lm_x, const double* __restrict__ lm_y,
                                                         uint32_t L, const double* __restrict__ poses,
                                                         const float* __restrict__ noise, uint32_t nscan, double sigma,
                                                         double radius, double rmin, double rmax,
                                                         float* __restrict__ ranges) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)nscan * RAYS) return;
    const uint32_t scan = (uint32_t)(idx / RAYS), k = (uint32_t)(idx % RAYS);
    const double px = poses[3 * (size_t)scan], py = poses[3 * (size_t)scan + 1], yaw = poses[3 * (size_t)scan + 2];
    double sn, cs;
    sincos(yaw + (double)k * (3.14159265358979323846 / 180.0), &sn, &cs);
    double best = INFINITY;
    for (uint32_t l = 0; l < L; ++l) {
        const double cx = lm_x[l] - px, cy = lm_y[l] - py;
        const double b = cs * cx + sn * cy;
        const double disc = b * b - (cx * cx + cy * cy - radius * radius);
        if (disc >= 0.0) {
            const double t = b - sqrt(disc);
            if (t > 0.0 && t < best) best = t;
        }
    }
    float out = INFINITY;
    if (best >= rmin && best <= rmax) out = (float)(best + sigma * (noise ? (double)noise[idx] : 0.0));
    ranges[idx] = out;
}

}  // namespace

}  // namespace aslam

using namespace aslam;

struct aslam_laser {
    int device = 0;
    cudaStream_t stream = nullptr;
    uint32_t max_scans = 0;
    LaserParams lp;
    float* d_cos = nullptr;
    float* d_sin = nullptr;
    float* d_ranges = nullptr;   // staging for the host entry point
    float* d_range_out = nullptr;
    float* d_bearing_out = nullptr;
    int32_t* d_counts = nullptr;
    cudaStream_t copy_in = nullptr;   // host path: the scans of chunk k+1 cross PCIe while chunk k is processed
    cudaEvent_t ev_in = nullptr;
    uint2* d_work = nullptr;          // accepted clusters of one batch: (output slot, start | length << 16)
    uint32_t* d_work_count = nullptr;
    uint64_t launches = 0;
};

extern "C" {

void aslam_laser_config_default(aslam_laser_config* c) {
    if (!c) return;
    c->deg2rad = 0.01745329251f;  // config.h:40
    c->min_dist_thresh = 0.5f;    // config.h:43
    c->std_max = 0.4f;            // config.h:48
    c->min_mean = 1.5f;           // config.h:49
    c->max_mean = 3.0f;           // config.h:50
    c->min_cluster_points = 3;    // config.h:51
}

aslam_status aslam_laser_create(const aslam_laser_config* cfg, uint32_t max_scans, int32_t device, aslam_laser** out) {
    if (!out || max_scans == 0) return ASLAM_ERR_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("laser create: no CUDA device (this library has no CPU path)");
        return ASLAM_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) return ASLAM_ERR_ARG;
    aslam_laser_config c;
    if (cfg)
        c = *cfg;
    else
        aslam_laser_config_default(&c);
    ASLAM_CUDA(cudaSetDevice(device));
    aslam_laser* h = new (std::nothrow) aslam_laser;
    if (!h) return ASLAM_ERR_ARG;
    h->device = device;
    h->max_scans = max_scans;
    h->lp.thresh = c.min_dist_thresh;
    h->lp.std_max = c.std_max;
    h->lp.min_mean = c.min_mean;
    h->lp.max_mean = c.max_mean;
    h->lp.min_pts = c.min_cluster_points;
    // LandMarks::initialize (sensor_landmark.cpp:50-57): float argument, float sin / cos of the host libm.
    // The clustering loop calls `bearing2pose(theta, msg->ranges[theta++])` (:87,98): the evaluation order of the two
    // arguments is unspecified and GCC -- the compiler of every ROS build; oracle/_ref, the reference compiled here,
    // pins it -- takes them right to left, so ray t is projected with table entry t + 1 (every cluster, and every
    // published bearing, is rotated by +1 degree) and ray 359 with entry 360, one past the table, which reads 0.0f
    // from the vector's spare capacity. The device tables hold that per-ray entry directly.
    float hc[RAYS], hs[RAYS];
    for (int th = 0; th < RAYS; ++th) {
        hs[th] = th + 1 < RAYS ? std::sin(c.deg2rad * static_cast<float>(th + 1)) : 0.0f;
        hc[th] = th + 1 < RAYS ? std::cos(c.deg2rad * static_cast<float>(th + 1)) : 0.0f;
    }
    h->lp.c0 = std::cos(c.deg2rad * static_cast<float>(0));
    h->lp.s0 = std::sin(c.deg2rad * static_cast<float>(0));
    h->lp.c359 = std::cos(c.deg2rad * static_cast<float>(RAYS - 1));
    h->lp.s359 = std::sin(c.deg2rad * static_cast<float>(RAYS - 1));
    bool ok = cuda_ok(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking), "cudaStreamCreate");
    ok = ok && cuda_ok(cudaStreamCreateWithFlags(&h->copy_in, cudaStreamNonBlocking), "cudaStreamCreate");
    ok = ok && cuda_ok(cudaEventCreateWithFlags(&h->ev_in, cudaEventDisableTiming), "cudaEventCreate");
    const size_t ob = (size_t)max_scans * ASLAM_LASER_MAX_LANDMARKS * sizeof(float);
    ok = ok && cuda_ok(cudaMalloc((void**)&h->d_cos, sizeof(hc)), "cudaMalloc");
    ok = ok && cuda_ok(cudaMalloc((void**)&h->d_sin, sizeof(hs)), "cudaMalloc");
    ok = ok && cuda_ok(cudaMalloc((void**)&h->d_ranges, (size_t)max_scans * RAYS * sizeof(float)), "cudaMalloc");
    ok = ok && cuda_ok(cudaMalloc((void**)&h->d_range_out, ob), "cudaMalloc");
    ok = ok && cuda_ok(cudaMalloc((void**)&h->d_bearing_out, ob), "cudaMalloc");
    ok = ok && cuda_ok(cudaMalloc((void**)&h->d_counts, (size_t)max_scans * sizeof(int32_t)), "cudaMalloc");
    ok = ok && cuda_ok(cudaMalloc((void**)&h->d_work, (size_t)max_scans * ASLAM_LASER_MAX_LANDMARKS * sizeof(uint2)), "cudaMalloc");
    ok = ok && cuda_ok(cudaMalloc((void**)&h->d_work_count, sizeof(uint32_t)), "cudaMalloc");
    ok = ok && cuda_ok(cudaMemcpy(h->d_cos, hc, sizeof(hc), cudaMemcpyHostToDevice), "tables");
    ok = ok && cuda_ok(cudaMemcpy(h->d_sin, hs, sizeof(hs), cudaMemcpyHostToDevice), "tables");
    if (!ok) {
        aslam_laser_destroy(h);
        return ASLAM_ERR_CUDA;
    }
    *out = h;
    return ASLAM_OK;
}

aslam_status aslam_laser_destroy(aslam_laser* h) {
    if (!h) return ASLAM_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    void* ptrs[] = {h->d_cos, h->d_sin, h->d_ranges, h->d_range_out, h->d_bearing_out, h->d_counts, h->d_work, h->d_work_count};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    if (h->copy_in) {
        cudaStreamSynchronize(h->copy_in);
        cudaStreamDestroy(h->copy_in);
    }
    if (h->ev_in) cudaEventDestroy(h->ev_in);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return ASLAM_OK;
}

aslam_status aslam_laser_extract_device(aslam_laser* h, const float* d_ranges, uint32_t nscan, float* d_range_out,
                                        float* d_bearing_out, int32_t* d_counts) {
    ASLAM_NVTX("aslam_laser_extract_device");
    if (!h || !d_ranges || !d_range_out || !d_bearing_out || !d_counts) return ASLAM_ERR_ARG;
    if (nscan == 0) return ASLAM_OK;
    ASLAM_CUDA(cudaSetDevice(h->device));
    // the work list holds the accepted clusters of at most max_scans scans: larger batches go in chunks
    for (uint32_t first = 0; first < nscan; first += h->max_scans) {
        const uint32_t cnt = nscan - first < h->max_scans ? nscan - first : h->max_scans;
        const size_t o = (size_t)first * ASLAM_LASER_MAX_LANDMARKS;
        ASLAM_CUDA(cudaMemsetAsync(h->d_work_count, 0, sizeof(uint32_t), h->stream));
        laser_extract_kernel<<<(cnt + LWPB - 1) / LWPB, LWPB * 32, 0, h->stream>>>(
            d_ranges + (size_t)first * RAYS, cnt, h->d_cos, h->d_sin, h->lp, h->d_work, h->d_work_count, d_counts + first);
        // worst-case grid (every scan with a full list); CTAs beyond the work count return at once
        const uint32_t fit_grid = (cnt * ASLAM_LASER_MAX_LANDMARKS + 127) / 128;
        laser_fit_kernel<<<fit_grid, 128, 0, h->stream>>>(d_ranges + (size_t)first * RAYS, h->d_cos, h->d_sin, h->d_work,
                                                         h->d_work_count, d_range_out + o, d_bearing_out + o);
        ASLAM_CUDA(cudaGetLastError());
        h->launches += 2;
    }
    return ASLAM_OK;
}

aslam_status aslam_laser_extract(aslam_laser* h, const float* ranges, uint32_t nscan, float* range_out,
                                 float* bearing_out, int32_t* counts) {
    ASLAM_NVTX("aslam_laser_extract");
    if (!h || !ranges || !range_out || !bearing_out || !counts || nscan > h->max_scans) return ASLAM_ERR_ARG;
    if (nscan == 0) return ASLAM_OK;
    ASLAM_CUDA(cudaSetDevice(h->device));
    // Chunks of 8,192 scans: the host->device copy of a chunk runs on its own stream, the kernels and the (much smaller)
    // result copies of the previous chunk on the handle's stream -- with pinned host buffers the call is bound by the
    // input crossing PCIe once (11.8 MB per chunk), not by the sum of copy + kernels + copy.
    constexpr uint32_t CHUNK = 8192;
    // the staging area must not be overwritten while an earlier call's kernels still read it
    ASLAM_CUDA(cudaEventRecord(h->ev_in, h->stream));
    ASLAM_CUDA(cudaStreamWaitEvent(h->copy_in, h->ev_in, 0));
    for (uint32_t first = 0; first < nscan; first += CHUNK) {
        const uint32_t cnt = nscan - first < CHUNK ? nscan - first : CHUNK;
        const size_t o = (size_t)first * ASLAM_LASER_MAX_LANDMARKS, ob = (size_t)cnt * ASLAM_LASER_MAX_LANDMARKS * sizeof(float);
        ASLAM_CUDA(cudaMemcpyAsync(h->d_ranges + (size_t)first * RAYS, ranges + (size_t)first * RAYS,
                                   (size_t)cnt * RAYS * sizeof(float), cudaMemcpyHostToDevice, h->copy_in));
        ASLAM_CUDA(cudaEventRecord(h->ev_in, h->copy_in));
        ASLAM_CUDA(cudaStreamWaitEvent(h->stream, h->ev_in, 0));
        const aslam_status rc = aslam_laser_extract_device(h, h->d_ranges + (size_t)first * RAYS, cnt, h->d_range_out + o,
                                                           h->d_bearing_out + o, h->d_counts + first);
        if (rc != ASLAM_OK) return rc;
        ASLAM_CUDA(cudaMemcpyAsync(range_out + o, h->d_range_out + o, ob, cudaMemcpyDeviceToHost, h->stream));
        ASLAM_CUDA(cudaMemcpyAsync(bearing_out + o, h->d_bearing_out + o, ob, cudaMemcpyDeviceToHost, h->stream));
        ASLAM_CUDA(cudaMemcpyAsync(counts + first, h->d_counts + first, (size_t)cnt * sizeof(int32_t), cudaMemcpyDeviceToHost,
                                   h->stream));
    }
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}

aslam_status aslam_laser_synthesize_device(aslam_laser* h, const double* d_lm_x, const double* d_lm_y, uint32_t L,
                                           const double* d_poses, const float* d_noise, uint32_t nscan, double sigma,
                                           double radius, float* d_ranges) {
    if (!h || !d_lm_x || !d_lm_y || !d_poses || !d_ranges) return ASLAM_ERR_ARG;
    if (nscan == 0) return ASLAM_OK;
    ASLAM_CUDA(cudaSetDevice(h->device));
    const size_t total = (size_t)nscan * RAYS;
    laser_synth_kernel<<<(unsigned)((total + 255) / 256), 256, 0, h->stream>>>(d_lm_x, d_lm_y, L, d_poses, d_noise, nscan,
                                                                              sigma, radius, 0.12, 3.5, d_ranges);
    ASLAM_CUDA(cudaGetLastError());
    h->launches += 1;
    return ASLAM_OK;
}

aslam_status aslam_laser_synthesize(aslam_laser* h, const double* lm_x, const double* lm_y, uint32_t L, const double* poses,
                                    const float* noise, uint32_t nscan, double sigma, double radius, float* ranges) {
    if (!h || !lm_x || !lm_y || !poses || !ranges || nscan > h->max_scans) return ASLAM_ERR_ARG;
    if (nscan == 0) return ASLAM_OK;
    ASLAM_CUDA(cudaSetDevice(h->device));
    double* d_map = nullptr;   // [2][L] map, then [nscan][3] poses
    float* d_noise = nullptr;
    const size_t total = (size_t)nscan * RAYS;
    ASLAM_CUDA(cudaMalloc((void**)&d_map, ((size_t)2 * L + (size_t)3 * nscan) * sizeof(double)));
    aslam_status rc = ASLAM_OK;
    if (noise && !cuda_ok(cudaMalloc((void**)&d_noise, total * sizeof(float)), "cudaMalloc")) rc = ASLAM_ERR_CUDA;
    const cudaMemcpyKind k = cudaMemcpyHostToDevice;
    if (rc == ASLAM_OK && !cuda_ok(cudaMemcpyAsync(d_map, lm_x, L * sizeof(double), k, h->stream), "copy")) rc = ASLAM_ERR_CUDA;
    if (rc == ASLAM_OK && !cuda_ok(cudaMemcpyAsync(d_map + L, lm_y, L * sizeof(double), k, h->stream), "copy")) rc = ASLAM_ERR_CUDA;
    if (rc == ASLAM_OK && !cuda_ok(cudaMemcpyAsync(d_map + 2 * L, poses, (size_t)3 * nscan * sizeof(double), k, h->stream), "copy"))
        rc = ASLAM_ERR_CUDA;
    if (rc == ASLAM_OK && noise && !cuda_ok(cudaMemcpyAsync(d_noise, noise, total * sizeof(float), k, h->stream), "copy"))
        rc = ASLAM_ERR_CUDA;
    if (rc == ASLAM_OK)
        rc = aslam_laser_synthesize_device(h, d_map, d_map + L, L, d_map + 2 * L, d_noise, nscan, sigma, radius, h->d_ranges);
    if (rc == ASLAM_OK && !cuda_ok(cudaMemcpyAsync(ranges, h->d_ranges, total * sizeof(float), cudaMemcpyDeviceToHost, h->stream), "copy"))
        rc = ASLAM_ERR_CUDA;
    if (!cuda_ok(cudaStreamSynchronize(h->stream), "sync") && rc == ASLAM_OK) rc = ASLAM_ERR_CUDA;
    cudaFree(d_map);
    if (d_noise) cudaFree(d_noise);
    return rc;
}

aslam_status aslam_laser_sync(aslam_laser* h) {
    if (!h) return ASLAM_ERR_ARG;
    ASLAM_CUDA(cudaStreamSynchronize(h->stream));
    return ASLAM_OK;
}
void* aslam_laser_stream(aslam_laser* h) { return h ? (void*)h->stream : nullptr; }
uint64_t aslam_laser_launch_count(aslam_laser* h) { return h ? h->launches : 0; }

}  // extern "C"
