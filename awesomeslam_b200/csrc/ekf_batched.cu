// K1: batched EKF-SLAM step. Replaces, for every filter of a batch,
// updateZandA + slam of the reference (src/ekf/ekf.cpp:137-213, 293-311) and the models of
// include/awesome_slam/common.h:46-90.
//
// Mapping. A filter is served by TPF = NP/CPT threads; thread t owns CPT adjacent COLUMNS of
// the covariance and keeps them in registers (NP doubles per column) for the whole launch --
// one step, or `steps` replayed messages with P never leaving the chip. A CTA holds FPB
// filters (about 128 threads); X, Z and the few broadcast vectors live in shared memory.
//
// Math. The reference forms the dense N x N S = H P H' + R and K = P H' S^-1 (18 N^3 flop).
// R is diagonal and H is linearised once per step, so the joint update equals a sequence of
// block updates with that FIXED linearisation: the 3 odometry rows (H = I), then the 2 rows
// (range, bearing) of each landmark, with the innovation of a block taken against the
// running state, y_g = Y0_g - H_g (x - x_pred). That is N^3 + O(N^2) fma instead of 9 N^3
// and needs no N x N factorisation. For block g with rows H_g (5 non-zeros per row):
//   V = H_g P            thread j: v_a[j] = sum_c H_g[a][c] P[c][j]     (own column only)
//   S_g = V H_g' + R_g   2x2 (3x3 for the pose block), every thread redundantly from smem
//   G = S_g^-1 V         own column
//   P[i][j] -= sum_a V[a][i] G[a][j]    V[.][i] broadcast from smem, 2 (3) fma per element
//   dx[j]  += sum_a G[a][j] y_a
// (P H_g' is taken as (H_g P)' -- P is symmetric to rounding in the reference as well,
// SURVEY.md section 6.2.) One __syncthreads per block; V and dx are double-buffered.
#include <cuda_runtime.h>

#include "ekf_batched_common.cuh"

namespace aslam {

namespace {

template <int NP, int CPT>
struct Geo {
    static constexpr int TPF = NP / CPT;
    static constexpr int FPB = (128 / TPF) > 0 ? (128 / TPF) : 1;
    static constexpr int THREADS = TPF * FPB;
    static constexpr int LMAX = (NP - 3) / 2;
    // N <= max_state_dim - 1 <= NP - 1 (batched_padded_dim): row NP - 1 of every covariance is padding (zero), so the
    // register tile holds NR = NP - 1 rows per column
    static constexpr int NR = NP - 1;
    static_assert(NP % CPT == 0, "NP must be a multiple of CPT");
};

template <int NP>
struct __align__(16) EkfSlot {
    double V[2][NP * 4];  // block rows of H_g P: pitch 4 for the pose block, 2 for landmark blocks
    double dl[2][NP];     // running state correction x - x_pred
    double X[NP];
    double Z[NP];
    double Y0[NP];        // Z - h(x_pred), bearings wrapped (ekf.cpp:302-307)
    double H[(NP - 3) / 2 + 1][4];
    double a00, a10;
    float Xf[NP];         // float copy of X for the association distances (ekf.cpp:161,166 round to float)
    MsgBuf msg[2];        // this step's message and the next one (in flight)
    FrontSlot fr;
};

template <int NP, int CPT>
__global__ void __launch_bounds__(Geo<NP, CPT>::THREADS) ekf_batched_kernel(const EkfParams kp) {
    using G = Geo<NP, CPT>;
    constexpr int TPF = G::TPF;
    constexpr int NR = G::NR;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    EkfSlot<NP>* slots = reinterpret_cast<EkfSlot<NP>*>(smem_raw);
    const int fs = threadIdx.x / TPF, ct = threadIdx.x % TPF;
    const uint32_t f = blockIdx.x * G::FPB + fs;
    const bool live = f < kp.st.batch;
    EkfSlot<NP>& s = slots[fs];
    const int j0 = ct * CPT;  // first owned column
    __shared__ int s_lcta;    // landmark blocks any filter of this CTA needs this step

    double p[CPT][NR];
    {
        const double* Pg = kp.st.P + (size_t)(live ? f : 0) * NP * NP;
#pragma unroll
        for (int i = 0; i < NR; ++i) {
            if constexpr (CPT == 2) {
                const double2 v = live ? *reinterpret_cast<const double2*>(Pg + i * NP + j0) : make_double2(0, 0);
                p[0][i] = v.x;
                p[1][i] = v.y;
            } else {
#pragma unroll
                for (int c = 0; c < CPT; ++c) p[c][i] = live ? Pg[i * NP + j0 + c] : 0.0;
            }
        }
    }
    for (int j = ct; j < NP; j += TPF) {
        s.X[j] = live ? kp.st.X[(size_t)f * NP + j] : 0.0;
        s.Z[j] = live ? kp.st.Z[(size_t)f * NP + j] : 0.0;
        s.Xf[j] = (float)s.X[j];
    }
    if (ct == 0) {
        s.fr.N = live ? kp.st.N[f] : 3;
        s.fr.status = live ? kp.st.status[f] : 0;
        s.fr.flags = live ? kp.st.flags[f] : 0;
        s.fr.wcount = live ? kp.st.wcount[f] : 0;
        s.fr.promote = 0;
        s.a00 = live ? kp.st.A[2 * (size_t)f] : 1.0;
        s.a10 = live ? kp.st.A[2 * (size_t)f + 1] : 0.0;
    }
    if (live) prefetch_message<TPF>(s.msg[0], kp.in, kp.st.batch, f, 0, ct);
    cp_async_wait_all();
    __syncthreads();

    for (uint32_t step = 0; step < kp.in.steps; ++step) {
        // ---- cbOdom gate (ekf.cpp:76): nothing happens before the first observation message
        const bool gated = !live || (s.fr.flags & FLAG_INIT_Z);
        if (threadIdx.x == 0) s_lcta = 0;
        // this step's message has landed (waited for behind the previous step); the next one starts now and is in
        // flight for the whole step
        const MsgBuf& m = s.msg[step & 1];
        if (live && step + 1 < kp.in.steps) prefetch_message<TPF>(s.msg[(step + 1) & 1], kp.in, kp.st.batch, f, step + 1, ct);
        // ---- updateZandA, ekf.cpp:139-142
        if (!gated) {
            if (ct == 0) {
                s.fr.ox = m.ox;
                s.fr.oy = m.oy;
                s.fr.yaw = m.yaw;
                s.fr.vx = m.vx;
                s.fr.az = m.az;
                s.fr.dt = m.dt;
                s.Z[0] = m.ox;
                s.Z[1] = m.oy;
                s.Z[2] = (double)m.yaw;
                uint32_t cnt = m.cnt;
                if (cnt > ASLAM_MAX_OBS) {
                    cnt = ASLAM_MAX_OBS;
                    s.fr.status |= ASLAM_ST_OBS_OVERFLOW;
                }
                s.fr.obs_base = m.base;
                s.fr.nobs = cnt;
                s.fr.unassoc = 0;
            }
            for (int l = ct; l < (int)(sizeof(s.fr.owner) / sizeof(int32_t)); l += TPF) s.fr.owner[l] = 0;
        }
        __syncthreads();
        if (!gated) front_phase_a<true>(s.fr, kp.fp, s.Xf, s.Z, m.rng, m.brg, 0u, f, ct, TPF);
        __syncthreads();
        if (!gated) {
            int32_t* assoc = nullptr;
            if (kp.in.assoc && step + 1 == kp.in.steps) assoc = kp.in.assoc + (kp.in.offsets ? s.fr.obs_base : f * kp.in.obs_per_filter);
            front_phase_b_assoc(s.fr, s.Z, assoc, ct, TPF);
        }
        if (ct == 0 && !gated) {
            front_phase_b_wait(s.fr, kp.fp, s.X, s.Z, f);
            // ekf.cpp:87-91 X = Z on the first processed message: the pose rows here, the landmark rows
            // (other threads may still be writing Z) behind the next barrier
            const bool initx = s.fr.flags & FLAG_INIT_X;
            // slam(): ekf.cpp:295-296
            double x = initx ? s.Z[0] : s.X[0], y = initx ? s.Z[1] : s.X[1], th = initx ? s.Z[2] : s.X[2];
            motion_model(x, y, th, (float)s.fr.vx, (float)s.fr.az, s.fr.dt);
            s.X[0] = x;
            s.X[1] = y;
            s.X[2] = (double)normalize_angle((float)th, kp.fp.pi);
            atomicMax(&s_lcta, (int)(s.fr.N - 3) / 2);
        }
        // ekf.cpp:206-212 on the filter's second thread, beside the pose predict of the first: the two are independent
        // (inputs: the message and Z[2], the odometry yaw) and each is a chain of two double-precision sincos
        if (ct == 1 && !gated) motion_jacobian(s.fr.vx, s.fr.az, s.fr.dt, s.Z[2], s.a00, s.a10);
        __syncthreads();
        const uint32_t N = s.fr.N;
        const int lcta = s_lcta;
        const bool initx = !gated && (s.fr.flags & FLAG_INIT_X);
        if (initx)
            for (uint32_t j = 3 + ct; j < N; j += TPF) s.X[j] = s.Z[j];
        // ---- growth of P (ekf.cpp:277): new diagonal entries, the rest of the new rows/cols is 0
        if (s.fr.grew_from && !gated) {
            const uint32_t g0 = s.fr.grew_from;
            // (arithmetic mask, not `if (i == j) p[i] = ..`: the latter is turned into a dynamically
            // indexed store, which would move the whole register tile to local memory)
#pragma unroll
            for (int c = 0; c < CPT; ++c) {
                const uint32_t j = j0 + c;
                const double dj = (j >= g0 && j < N) ? kp.kp_lm : 0.0;
#pragma unroll
                for (int i = 0; i < NR; ++i) p[c][i] = fma((double)((uint32_t)i == j), dj, p[c][i]);
            }
        }
        // ---- P = A P A' + Q (ekf.cpp:297); A = I except A(0,0), A(1,0)
        if (!gated) {
            const double a00 = s.a00, a10 = s.a10;
#pragma unroll
            for (int c = 0; c < CPT; ++c) {
                const double p0 = p[c][0], p1 = p[c][1];
                p[c][0] = a00 * p0;
                p[c][1] = a10 * p0 + p1;
            }
            if (j0 == 0) {
#pragma unroll
                for (int i = 0; i < NR; ++i) s.V[0][i] = p[0][i];  // column 0 of A P
            }
        }
        __syncthreads();
        if (!gated) {
            const double a00 = s.a00, a10 = s.a10;
#pragma unroll
            for (int c = 0; c < CPT; ++c) {
                const int j = j0 + c;
                if (j == 0) {
#pragma unroll
                    for (int i = 0; i < NR; ++i) p[c][i] = p[c][i] * a00;
                } else if (j == 1) {
#pragma unroll
                    for (int i = 0; i < NR; ++i) p[c][i] = s.V[0][i] * a10 + p[c][i];
                }
#pragma unroll
                for (int i = 0; i < 3; ++i)
                    if (i == j) p[c][i] += kp.kq;
            }
            // ---- updateH (ekf.cpp:117-134) and Y = Z - h(X) (ekf.cpp:302-307), one landmark per thread
            const double x = s.X[0], y = s.X[1], th = s.X[2];
            for (uint32_t l = ct; 3 + 2 * l < N; l += TPF) {
                const double lx = s.X[3 + 2 * l], ly = s.X[4 + 2 * l];
                const HEntries h = jacobian_entries(x, y, lx, ly);
                s.H[l][0] = h.a;
                s.H[l][1] = h.b;
                s.H[l][2] = h.c;
                s.H[l][3] = h.d;
                double rg, br;
                measure_landmark(x, y, th, lx, ly, rg, br);
                s.Y0[3 + 2 * l] = s.Z[3 + 2 * l] - rg;
                s.Y0[4 + 2 * l] = (double)normalize_angle((float)(s.Z[4 + 2 * l] - br), kp.fp.pi);
            }
            if (ct == 0) {
                if (initx) s.fr.flags &= ~FLAG_INIT_X;
                s.Y0[0] = s.Z[0] - x;
                s.Y0[1] = s.Z[1] - y;
                s.Y0[2] = (double)normalize_angle((float)(s.Z[2] - th), kp.fp.pi);
            }
        }
        __syncthreads();  // column 0 of AP consumed; H, Y0 visible

        // ---- pose block: rows 0..2 of H are the identity (ekf.cpp:61,276)
        double dl[CPT];
        if (!gated) {
#pragma unroll
            for (int c = 0; c < CPT; ++c) {
                double* v = &s.V[0][(j0 + c) * 4];
                v[0] = p[c][0];
                v[1] = p[c][1];
                v[2] = p[c][2];
            }
        }
        __syncthreads();
        if (!gated) {
            double S[3][3];
#pragma unroll
            for (int a = 0; a < 3; ++a)
#pragma unroll
                for (int b = 0; b < 3; ++b) S[a][b] = s.V[0][b * 4 + a] + (a == b ? kp.kr_pose : 0.0);
            // 3x3 inverse by cofactors
            const double c00 = S[1][1] * S[2][2] - S[1][2] * S[2][1];
            const double c01 = S[1][2] * S[2][0] - S[1][0] * S[2][2];
            const double c02 = S[1][0] * S[2][1] - S[1][1] * S[2][0];
            const double det = S[0][0] * c00 + S[0][1] * c01 + S[0][2] * c02;
            const double id = fast_rcp(det);
            double Si[3][3];
            Si[0][0] = c00 * id;
            Si[1][0] = c01 * id;
            Si[2][0] = c02 * id;
            Si[0][1] = (S[0][2] * S[2][1] - S[0][1] * S[2][2]) * id;
            Si[1][1] = (S[0][0] * S[2][2] - S[0][2] * S[2][0]) * id;
            Si[2][1] = (S[0][1] * S[2][0] - S[0][0] * S[2][1]) * id;
            Si[0][2] = (S[0][1] * S[1][2] - S[0][2] * S[1][1]) * id;
            Si[1][2] = (S[0][2] * S[1][0] - S[0][0] * S[1][2]) * id;
            Si[2][2] = (S[0][0] * S[1][1] - S[0][1] * S[1][0]) * id;
            const double y0 = s.Y0[0], y1 = s.Y0[1], y2 = s.Y0[2];
            double g[CPT][3];
#pragma unroll
            for (int c = 0; c < CPT; ++c) {
#pragma unroll
                for (int a = 0; a < 3; ++a)
                    g[c][a] = fma(Si[a][0], p[c][0], fma(Si[a][1], p[c][1], Si[a][2] * p[c][2]));
                dl[c] = fma(g[c][0], y0, fma(g[c][1], y1, g[c][2] * y2));
            }
#pragma unroll
            for (int i = 0; i < NR; ++i) {
                const double2 v01 = *reinterpret_cast<const double2*>(&s.V[0][i * 4]);
                const double v2 = s.V[0][i * 4 + 2];
#pragma unroll
                for (int c = 0; c < CPT; ++c)
                    p[c][i] = fma(-v01.x, g[c][0], fma(-v01.y, g[c][1], fma(-v2, g[c][2], p[c][i])));
                if ((i & 7) == 7) asm volatile("" ::: "memory");
            }
        } else {
#pragma unroll
            for (int c = 0; c < CPT; ++c) dl[c] = 0.0;
        }

        // ---- landmark blocks, rows (3+2l, 4+2l)
        int buf = 0;
#pragma unroll
        for (int l = 0; l < G::LMAX; ++l) {
            const int r0 = 3 + 2 * l, r1 = 4 + 2 * l;
            buf ^= 1;
            const bool act = !gated && (uint32_t)r1 < N;
            double v[CPT][2];
            double ha = 0, hb = 0, hc = 0, hd = 0;
            if (act) {
                const double2 h01 = *reinterpret_cast<const double2*>(&s.H[l][0]);
                const double2 h23 = *reinterpret_cast<const double2*>(&s.H[l][2]);
                ha = h01.x;
                hb = h01.y;
                hc = h23.x;
                hd = h23.y;
#pragma unroll
                for (int c = 0; c < CPT; ++c) {
                    // range row [a, c, 0, .., -a, -c]; bearing row [b, d, -1, .., -b, -d]
                    v[c][0] = fma(ha, p[c][0] - p[c][r0], hc * (p[c][1] - p[c][r1]));
                    v[c][1] = fma(hb, p[c][0] - p[c][r0], fma(hd, p[c][1] - p[c][r1], -p[c][2]));
                    *reinterpret_cast<double2*>(&s.V[buf][(j0 + c) * 2]) = make_double2(v[c][0], v[c][1]);
                    s.dl[buf][j0 + c] = dl[c];
                }
            }
            if (l >= lcta) break;  // no filter of this CTA has landmark l (CTA-uniform)
            __syncthreads();
            if (act) {
                const double2 w0 = *reinterpret_cast<const double2*>(&s.V[buf][0]);
                const double2 w1 = *reinterpret_cast<const double2*>(&s.V[buf][2]);
                const double2 w2 = *reinterpret_cast<const double2*>(&s.V[buf][4]);
                const double2 wa = *reinterpret_cast<const double2*>(&s.V[buf][r0 * 2]);
                const double2 wb = *reinterpret_cast<const double2*>(&s.V[buf][r1 * 2]);
                // S_g = V H_g' + R_g
                const double S00 = fma(ha, w0.x - wa.x, hc * (w1.x - wb.x)) + kp.kr_lm;
                const double S01 = fma(hb, w0.x - wa.x, fma(hd, w1.x - wb.x, -w2.x));
                const double S10 = fma(ha, w0.y - wa.y, hc * (w1.y - wb.y));
                const double S11 = fma(hb, w0.y - wa.y, fma(hd, w1.y - wb.y, -w2.y)) + kp.kr_lm;
                const double id = fast_rcp(S00 * S11 - S01 * S10);
                const double i00 = S11 * id, i01 = -S01 * id, i10 = -S10 * id, i11 = S00 * id;
                // innovation against the running state: y = Y0_g - H_g dx
                const double d0 = s.dl[buf][0], d1 = s.dl[buf][1], d2 = s.dl[buf][2];
                const double da = s.dl[buf][r0], db = s.dl[buf][r1];
                const double y0 = s.Y0[r0] - fma(ha, d0 - da, hc * (d1 - db));
                const double y1 = s.Y0[r1] - fma(hb, d0 - da, fma(hd, d1 - db, -d2));
                double g[CPT][2];
#pragma unroll
                for (int c = 0; c < CPT; ++c) {
                    g[c][0] = fma(i00, v[c][0], i01 * v[c][1]);
                    g[c][1] = fma(i10, v[c][0], i11 * v[c][1]);
                    dl[c] = fma(g[c][0], y0, fma(g[c][1], y1, dl[c]));
                }
#pragma unroll
                for (int i = 0; i < NR; ++i) {
                    const double2 vi = *reinterpret_cast<const double2*>(&s.V[buf][i * 2]);
#pragma unroll
                    for (int c = 0; c < CPT; ++c) p[c][i] = fma(-vi.x, g[c][0], fma(-vi.y, g[c][1], p[c][i]));
                    // keep the scheduler from hoisting all NP broadcast loads ahead of the FMAs (spills)
                    if ((i & 7) == 7) asm volatile("" ::: "memory");
                }
            }
        }
        __syncthreads();
        // ---- X = X + K Y (ekf.cpp:309)
        if (!gated) {
#pragma unroll
            for (int c = 0; c < CPT; ++c)
                if ((uint32_t)(j0 + c) < N) {
                    const double xn = s.X[j0 + c] + dl[c];
                    s.X[j0 + c] = xn;
                    s.Xf[j0 + c] = (float)xn;
                }
        }
        cp_async_wait_all();  // every thread waits for its own copies of the next message; the barrier publishes them
        __syncthreads();
    }

    // ---- write back
    if (live) {
        double* Pg = kp.st.P + (size_t)f * NP * NP;
        bool bad = false;
        double acc = 0.0;  // NaN/Inf anywhere in the owned columns makes the sum non-finite
#pragma unroll
        for (int c = 0; c < CPT; ++c) Pg[(NP - 1) * NP + j0 + c] = 0.0;  // the padding row
#pragma unroll
        for (int i = 0; i < NR; ++i) {
#pragma unroll
            for (int c = 0; c < CPT; ++c) acc += p[c][i];
            if constexpr (CPT == 2) {
                *reinterpret_cast<double2*>(Pg + i * NP + j0) = make_double2(p[0][i], p[1][i]);
            } else {
#pragma unroll
                for (int c = 0; c < CPT; ++c) Pg[i * NP + j0 + c] = p[c][i];
            }
        }
        bad |= !isfinite(acc);
        for (int j = ct; j < NP; j += TPF) {
            kp.st.X[(size_t)f * NP + j] = s.X[j];
            kp.st.Z[(size_t)f * NP + j] = s.Z[j];
            bad |= !isfinite(s.X[j]);
        }
        if (bad) atomicOr(&kp.st.status[f], (uint32_t)ASLAM_ST_NONFINITE);
        if (ct == 0) {
            kp.st.N[f] = s.fr.N;
            if (s.fr.status) atomicOr(&kp.st.status[f], s.fr.status);
            kp.st.flags[f] = s.fr.flags;
            kp.st.wcount[f] = s.fr.wcount;
            kp.st.A[2 * (size_t)f] = s.a00;
            kp.st.A[2 * (size_t)f + 1] = s.a10;
        }
    }
}

template <int NP, int CPT>
aslam_status launch_t(const EkfParams& kp, cudaStream_t stream) {
    using G = Geo<NP, CPT>;
    const size_t smem = sizeof(EkfSlot<NP>) * G::FPB;
    ASLAM_DYN_SMEM((ekf_batched_kernel<NP, CPT>), smem);
    const uint32_t grid = (kp.st.batch + G::FPB - 1) / G::FPB;
    ekf_batched_kernel<NP, CPT><<<grid, G::THREADS, smem, stream>>>(kp);
    ASLAM_CUDA(cudaGetLastError());
    return ASLAM_OK;
}

}  // namespace

aslam_status launch_ekf_batched(const aslam_config& cfg, const BatchState& st, const BatchInput& in,
                                cudaStream_t stream) {
    EkfParams kp;
    kp.st = st;
    kp.in = in;
    kp.fp.pi = cfg.pi;
    kp.fp.thresh = cfg.min_dist_thresh;
    kp.fp.occ = cfg.min_landmark_occ;
    kp.fp.cap = cfg.max_state_dim;
    kp.fp.wrange = st.wrange;
    kp.fp.wbearing = st.wbearing;
    kp.fp.whits = st.whits;
    kp.kq = (double)cfg.ekf_kq;
    kp.kr_pose = (double)cfg.ekf_kr;
    kp.kr_lm = (double)cfg.ukf_kr;             // ekf.cpp:278 uses the UKF constant, as written
    kp.kp_lm = (double)cfg.ukf_kp_landmark_pose;  // ekf.cpp:277
    switch (st.ld) {
        case 30: return launch_t<30, 2>(kp, stream);
        case 36: return launch_t<36, 2>(kp, stream);
        case 48: return launch_t<48, 2>(kp, stream);
        case 64: return launch_t<64, 1>(kp, stream);
        default: set_error("ekf_batched: unsupported padded dim"); return ASLAM_ERR_UNSUPPORTED;
    }
}

}  // namespace aslam
