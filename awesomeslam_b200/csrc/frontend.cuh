// Data association, wait-list and landmark promotion shared by the batched EKF and UKF
// kernels: the front half of cbOdom (ekf.cpp:137-203 / ukf.cpp:113-179), the wait-list
// (ekf.cpp:217-253 / ukf.cpp:184-220) and the bookkeeping half of updateNewLandmark
// (ekf.cpp:255-268,280-289 / ukf.cpp:222-235,247-256).
//
// A filter is served by TPF consecutive threads of a CTA (they may straddle warps, so the
// phases are separated by __syncthreads(), which every thread of the CTA reaches).
//   phase A (parallel): each observation is normalised, projected with the odometry pose and
//                       matched against the mapped landmarks; wait-list entries are projected.
//   phase B (one thread): the order-dependent part -- Z writes, wait-list hits/appends,
//                       promotion -- which only touches floats already computed in phase A.
#pragma once
#include "device_math.cuh"

namespace aslam {

struct FrontSlot {
    float2 opt[ASLAM_MAX_OBS];   // observation projected to the world frame (structures.h:104)
    float orng[ASLAM_MAX_OBS];
    float obrg[ASLAM_MAX_OBS];   // normalised bearing (ekf.cpp:150)
    float omind[ASLAM_MAX_OBS];
    int32_t ocorr[ASLAM_MAX_OBS];
    float2 wpt[ASLAM_MAX_WAIT];  // wait-list entries projected with the CURRENT pose (ekf.cpp:228-234)
    uint32_t nobs, obs_base, wcount, N, status, flags, grew_from;
    int32_t promote;
    // TRACK mode of phase A: owner[l] = 1 + the LAST observation associated with landmark l (the one
    // whose Z write survives in the reference's sequential loop), unassoc = some observation missed
    int32_t owner[(64 - 3) / 2 + 2];
    uint32_t unassoc;
    uint8_t ohit[ASLAM_MAX_OBS];  // TRACK: the observation was associated (decided with the pre-promotion N)
    double ox, oy, vx, az;
    float yaw, dt;
};

struct FrontParams {
    float pi, thresh;
    int32_t occ, cap;
    float* wrange;
    float* wbearing;
    uint32_t* whits;
};

// The functions below are templates on the slot type: any struct with FrontSlot's member names will do (the batched
// UKF keeps the arrays in storage it shares with its sigma-point matrices and passes a struct of pointers).
// Phase A. X, Z: the filter's state vectors in shared memory (Z[0..2] already hold the
// odometry pose); X may be a float copy of the state (the reference rounds the landmark coordinates
// to float before the distance, ekf.cpp:161,166). lane/width: this thread's index in the filter's
// thread group; observation o is range[obs_first + o]. TRACK: also fill owner[] / unassoc (both zeroed
// by the caller before the call).
template <bool TRACK = false, typename XT = double, typename Slot = FrontSlot>
__device__ __forceinline__ void front_phase_a(Slot& s, const FrontParams& fp, const XT* X,
                                              const double* Z, const float* range, const float* bearing,
                                              uint32_t obs_first, uint32_t f, int lane, int width) {
    const double z0 = Z[0], z1 = Z[1], z2 = Z[2];
    const uint32_t N = s.N;
    for (uint32_t o = lane; o < s.nobs; o += width) {
        const float r = range[obs_first + o];
        const float b = normalize_angle(bearing[obs_first + o], fp.pi);
        const float2 pt = to_point(r, b, z0, z1, z2);
        int32_t corr = 0;
        float mind = 0.0f;
        if (N > 3) {  // ekf.cpp:160-173, strict '<' keeps the first minimum
            mind = euler_distance(pt.x, pt.y, (float)X[3], (float)X[4]);
            for (uint32_t j = 2; j < N - 3; j += 2) {
                const float d = euler_distance(pt.x, pt.y, (float)X[3 + j], (float)X[4 + j]);
                if (d < mind) {
                    corr = (int32_t)j;
                    mind = d;
                }
            }
        }
        s.orng[o] = r;
        s.obrg[o] = b;
        s.opt[o] = pt;
        s.ocorr[o] = corr;
        s.omind[o] = mind;
        if (TRACK) {
            const bool hit = N > 3 && mind < fp.thresh;  // ekf.cpp:175
            s.ohit[o] = hit;
            if (hit)
                atomicMax(&s.owner[corr >> 1], (int32_t)o + 1);
            else
                s.unassoc = 1u;
        }
    }
    const float* wr = fp.wrange + (size_t)f * ASLAM_MAX_WAIT;
    const float* wb = fp.wbearing + (size_t)f * ASLAM_MAX_WAIT;
    for (uint32_t w = lane; w < s.wcount; w += width) s.wpt[w] = to_point(wr[w], wb[w], z0, z1, z2);
}

// updateNewLandmarkWait (ekf.cpp:217-253) for observation o; runs on one thread.
template <typename Slot>
__device__ __forceinline__ void wait_list_update(Slot& s, const FrontParams& fp, uint32_t f, uint32_t o) {
    float* wr = fp.wrange + (size_t)f * ASLAM_MAX_WAIT;
    float* wb = fp.wbearing + (size_t)f * ASLAM_MAX_WAIT;
    uint32_t* wh = fp.whits + (size_t)f * ASLAM_MAX_WAIT;
    bool push = true;
    if (s.wcount > 0) {
        const float2 p2 = s.opt[o];
        uint32_t corr = 0;
        float mind = euler_distance(p2.x, p2.y, s.wpt[0].x, s.wpt[0].y);
        for (uint32_t i = 1; i < s.wcount; ++i) {
            const float d = euler_distance(p2.x, p2.y, s.wpt[i].x, s.wpt[i].y);
            if (d < mind) {
                corr = i;
                mind = d;
            }
        }
        if (mind < fp.thresh) {
            const uint32_t h = wh[corr] + 1;
            wh[corr] = h;
            if (h == (uint32_t)fp.occ) s.promote = 1;
            push = false;
        }
    }
    if (push) {
        if (s.wcount < ASLAM_MAX_WAIT) {
            const uint32_t w = s.wcount;
            wr[w] = s.orng[o];
            wb[w] = s.obrg[o];
            wh[w] = 1;
            s.wpt[w] = s.opt[o];  // same function of the same floats as toPoint on the new entry
            s.wcount = w + 1;
            if (fp.occ == 1) s.promote = 1;
        } else {
            s.status |= ASLAM_ST_WAIT_OVERFLOW;
        }
    }
}

// Promotion of wait-list entries seen exactly `occ` times and growth of X/Z (ekf.cpp:190-203,
// 255-290; the caller grows P). One thread per filter.
template <typename Slot>
__device__ __forceinline__ void front_promote(Slot& s, const FrontParams& fp, double* X, double* Z, uint32_t f) {
    if (!s.promote) return;
    s.promote = 0;
    float* wr = fp.wrange + (size_t)f * ASLAM_MAX_WAIT;
    float* wb = fp.wbearing + (size_t)f * ASLAM_MAX_WAIT;
    uint32_t* wh = fp.whits + (size_t)f * ASLAM_MAX_WAIT;
    uint32_t fresh = 0;
    for (uint32_t w = 0; w < s.wcount; ++w) fresh += (wh[w] == (uint32_t)fp.occ);
    if (!fresh) return;
    const uint32_t cacheN = s.N;
    const uint32_t newN = cacheN + 2 * fresh;
    const bool accept = newN < (uint32_t)fp.cap;  // ekf.cpp:263
    uint32_t k = cacheN;
    for (uint32_t w = 0; w < s.wcount; ++w) {
        if (wh[w] != (uint32_t)fp.occ) continue;
        wh[w] += 1;  // ekf.cpp:195: bumped even when the batch is then rejected
        if (!accept) continue;
        const double r = (double)wr[w], b = (double)wb[w];
        Z[k] = r;
        Z[k + 1] = b;
        double sn, cs;
        sincos(__dadd_rn(Z[2], b), &sn, &cs);  // ekf.cpp:287-288
        X[k] = __dadd_rn(Z[0], __dmul_rn(r, cs));
        X[k + 1] = __dadd_rn(Z[1], __dmul_rn(r, sn));
        k += 2;
    }
    if (accept) {
        s.grew_from = cacheN;
        s.N = newN;
    } else {
        s.status |= ASLAM_ST_CAP_REJECTED;
    }
}

// Phase B, one thread per filter. Writes Z for associated observations, maintains the
// wait-list, promotes entries seen exactly `occ` times and grows X/Z (the caller grows P).
// assoc: per-observation output for this filter or nullptr.
template <typename Slot>
__device__ __forceinline__ void front_phase_b(Slot& s, const FrontParams& fp, double* X, double* Z,
                                              uint32_t f, int32_t* assoc) {
    s.grew_from = 0;
    for (uint32_t o = 0; o < s.nobs; ++o) {
        if (s.N > 3 && s.omind[o] < fp.thresh) {  // ekf.cpp:175-181
            Z[3 + s.ocorr[o]] = (double)s.orng[o];
            Z[4 + s.ocorr[o]] = (double)s.obrg[o];
            if (assoc) assoc[o] = s.ocorr[o];
            continue;
        }
        if (assoc) assoc[o] = -1;
        wait_list_update(s, fp, f, o);
    }
    front_promote(s, fp, X, Z, f);
}

// Phase B split for a TRACK-mode phase A. The Z writes of the associated observations are independent
// of the wait-list (which only reads floats computed in phase A), so they run on all the filter's
// threads: observation o writes iff it is the last one associated with its landmark -- exactly what
// survives the reference's sequential loop (ekf.cpp:175-181).
__device__ __forceinline__ void front_phase_b_assoc(const FrontSlot& s, double* Z,
                                                    int32_t* assoc, int lane, int width) {
    for (uint32_t o = lane; o < s.nobs; o += width) {
        const bool hit = s.ohit[o];  // not re-derived from s.N: the serial half may be growing it right now
        const int32_t corr = s.ocorr[o];
        if (hit && s.owner[corr >> 1] == (int32_t)o + 1) {
            Z[3 + corr] = (double)s.orng[o];
            Z[4 + corr] = (double)s.obrg[o];
        }
        if (assoc) assoc[o] = hit ? corr : -1;
    }
}
// ... and the order-dependent rest on one thread: wait-list hits/appends in observation order, promotion.
__device__ __forceinline__ void front_phase_b_wait(FrontSlot& s, const FrontParams& fp, double* X, double* Z,
                                                   uint32_t f) {
    s.grew_from = 0;
    if (s.unassoc) {
        for (uint32_t o = 0; o < s.nobs; ++o)
            if (!s.ohit[o]) wait_list_update(s, fp, f, o);
    }
    front_promote(s, fp, X, Z, f);
}

}  // namespace aslam
