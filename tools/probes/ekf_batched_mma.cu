// NOT PART OF THE PRODUCT LIBRARY (tools/probes/): a measured negative result kept for the record -- 74 M updates/s
// against 107 M for the kernel in csrc/ekf_batched.cu. It needs csrc/ekf_batched_common.cuh and
// a `case 36` hook in launch_ekf_batched to be built again.
//
// K1b: batched EKF-SLAM step, one WARP per filter, covariance update on the FP64 tensor pipe
// (mma.sync.m8n8k4.f64 = SASS DMMA.8x8x4). Same contract and same mathematics as ekf_batched.cu
// (updateZandA + slam, src/ekf/ekf.cpp:137-213, 293-311; models of include/awesome_slam/common.h:46-90),
// specialised for the cfg3 size: padded dimension 36 (up to 16 landmarks, N <= 35).
//
// Layout. The filter's P lives in the warp's registers for the whole launch as a 40 x 40 matrix of 5 x 5
// accumulator tiles (C fragments of the 8x8x4 MMA: lane (lr, lc) = (lane/4, lane%4) holds row lr and columns
// 2 lc, 2 lc + 1 of every tile): 50 doubles per lane. Rows are PERMUTED -- landmarks first (internal 0..31),
// pose last (32, 33, 34) -- so that the four state rows of a landmark pair never straddle a tile row.
// Internal column 35 is padding in P; it is used as the accumulator of the state correction x - x_pred.
//
// Math. Fixed-linearisation sequential block update (see ekf_batched.cu), in blocks of FOUR measurement rows: the
// pose block (3 rows, H = I), then one block per landmark pair (range, bearing, range, bearing). Per block:
//   support rows of P (3 pose rows + 4 landmark rows) -> shared memory     (10 predicated 16-byte stores per lane)
//   S = H_g P H_g' + R_g (4 x 4) from the 7 x 7 support sub-block, S = L L'  (Cholesky, every lane)
//   W = L^-1 H_g P (4 x 40): lane (lr, lc) computes W[lc][8t + lr], t = 0..4 -- exactly its A and B operands
//   P -= W' W: 25 DMMA; the B operand of column 35 is -L^-1 y, so the same DMMAs accumulate dx += W' L^-1 y
// There is no CTA barrier anywhere: a filter is a warp, phases are separated by __syncwarp().
#include <cuda_runtime.h>

#include "ekf_batched_common.cuh"

namespace aslam {

namespace {

constexpr int NPG = 36;    // padded dimension of the HBM layout (BatchState::ld)
constexpr int NT = 5;      // tiles per side
constexpr int NI = 8 * NT; // internal dimension
constexpr int LM = 16;     // landmark slots
constexpr int POSE = 32;   // internal index of pose row 0
constexpr int DXCOL = 35;  // internal column accumulating the state correction
constexpr int WPB = 4;     // warps (filters) per CTA

// internal index (< 35) -> reference state index
__device__ __forceinline__ int to_ref(int i) { return i < POSE ? i + 3 : i - POSE; }

struct __align__(16) MmaSlot {
    double Praw[7][NI];  // support rows of P: [0..2] pose, [3..6] the block's landmark rows
    double Hd[4][8];     // the block's 4 rows of H over the 7 support columns ([.][7] = 0)
    double T[4][8];      // H_g Psub (4 x 7); T[k][7] = innovation y_g[k]
    double S[4][4];
    double X[NPG], Z[NPG], Y0[NPG];
    double H[LM][4];
    double a00, a10;
    float Xf[NPG];
    MsgBuf msg[2];
    FrontSlot fr;
};

__device__ __forceinline__ void dmma(double2& c, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c.x), "+d"(c.y)
                 : "d"(a), "d"(b));
}

// One block of four measurement rows. TR: tile row holding the block's landmark rows (ignored for the pose block).
// b: landmark pair index (landmarks 2b, 2b+1), or -1 for the pose block.
template <int TR>
__device__ __forceinline__ void block_update(double2 (&c)[NT][NT], MmaSlot& s, const EkfParams& kp, int b, uint32_t N,
                                             int lane) {
    const int lr = lane >> 2, lc = lane & 3;
    const bool pose = b < 0;
    // ---- (1) support rows -> shared memory; dense H rows of the block
    if (lr < 3) {
#pragma unroll
        for (int tc = 0; tc < NT; ++tc) *reinterpret_cast<double2*>(&s.Praw[lr][8 * tc + 2 * lc]) = c[NT - 1][tc];
    }
    if (!pose && (lr >> 2) == (b & 1)) {
#pragma unroll
        for (int tc = 0; tc < NT; ++tc) *reinterpret_cast<double2*>(&s.Praw[3 + (lr & 3)][8 * tc + 2 * lc]) = c[TR][tc];
    }
    if (lr == 0) {  // lane lc writes H row k = lc
        const int k = lc;
        double h[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (pose) {
            if (k < 3) h[k] = 1.0;
        } else {
            const int l = 2 * b + (k >> 1);
            if ((uint32_t)(4 + 2 * l) < N) {
                const double2 h01 = *reinterpret_cast<const double2*>(&s.H[l][0]);
                const double2 h23 = *reinterpret_cast<const double2*>(&s.H[l][2]);
                // range row [a, c, 0, .., -a, -c]; bearing row [b, d, -1, .., -b, -d] (ekf.cpp:124-132)
                const double p0 = (k & 1) ? h01.y : h01.x, p1 = (k & 1) ? h23.y : h23.x;
                h[0] = p0;
                h[1] = p1;
                h[2] = (k & 1) ? -1.0 : 0.0;
                h[3 + 2 * (k >> 1)] = -p0;
                h[4 + 2 * (k >> 1)] = -p1;
            }
        }
#pragma unroll
        for (int e = 0; e < 8; e += 2) *reinterpret_cast<double2*>(&s.Hd[k][e]) = make_double2(h[e], h[e + 1]);
    }
    __syncwarp();
    // ---- (2) T = H_g Psub: lane (k = lc, c' = lr < 7); lane lr == 7: innovation against the running state
    const int nsup = pose ? 3 : 7;
    {
        double hk[7];
#pragma unroll
        for (int e = 0; e < 7; ++e) hk[e] = s.Hd[lc][e];
        const int col = lr < 3 ? POSE + lr : (pose ? DXCOL : 4 * b + (lr - 3));  // lr == 7 reads column 35 = dx
        const int colr = lr == 7 ? DXCOL : col;
        double t = 0.0;
#pragma unroll
        for (int e = 0; e < 7; ++e)
            if (e < nsup) t = fma(hk[e], s.Praw[e][colr], t);
        if (lr == 7) {
            double y0 = 0.0;
            if (pose) {
                if (lc < 3) y0 = s.Y0[lc];
            } else {
                const int l = 2 * b + (lc >> 1);
                if ((uint32_t)(4 + 2 * l) < N) y0 = s.Y0[3 + 2 * l + (lc & 1)];
            }
            t = y0 - t;
        } else if (lr >= nsup) {
            t = 0.0;
        }
        s.T[lc][lr] = t;
    }
    __syncwarp();
    // ---- (3) S = T H_g' + R: lanes lr < 4 compute S[k = lc][k' = lr]
    if (lr < 4) {
        double sum = 0.0;
#pragma unroll
        for (int e = 0; e < 7; ++e)
            if (e < nsup) sum = fma(s.T[lc][e], s.Hd[lr][e], sum);
        if (lc == lr) sum += pose ? (lc < 3 ? kp.kr_pose : 1.0) : kp.kr_lm;
        s.S[lc][lr] = sum;
    }
    __syncwarp();
    // ---- (4) Cholesky S = L L' and M = L^-1 (lower), every lane
    double m[4];  // row k = lc of M
    {
        const double2 r0 = *reinterpret_cast<const double2*>(&s.S[0][0]);
        const double2 r1 = *reinterpret_cast<const double2*>(&s.S[1][0]);
        const double2 r2a = *reinterpret_cast<const double2*>(&s.S[2][0]);
        const double2 r2b = *reinterpret_cast<const double2*>(&s.S[2][2]);
        const double2 r3a = *reinterpret_cast<const double2*>(&s.S[3][0]);
        const double2 r3b = *reinterpret_cast<const double2*>(&s.S[3][2]);
        const double i0 = rsqrt(r0.x);
        const double l10 = r1.x * i0, l20 = r2a.x * i0, l30 = r3a.x * i0;
        const double i1 = rsqrt(fma(-l10, l10, r1.y));
        const double l21 = fma(-l20, l10, r2a.y) * i1, l31 = fma(-l30, l10, r3a.y) * i1;
        const double i2 = rsqrt(fma(-l21, l21, fma(-l20, l20, r2b.x)));
        const double l32 = fma(-l31, l21, fma(-l30, l20, r3b.x)) * i2;
        const double i3 = rsqrt(fma(-l32, l32, fma(-l31, l31, fma(-l30, l30, r3b.y))));
        const double m10 = -(l10 * i0) * i1;
        const double m20 = -fma(l21, m10, l20 * i0) * i2, m21 = -(l21 * i1) * i2;
        const double m30 = -fma(l32, m20, fma(l31, m10, l30 * i0)) * i3, m31 = -fma(l32, m21, l31 * i1) * i3;
        const double m32 = -(l32 * i2) * i3;
        m[0] = lc == 0 ? i0 : lc == 1 ? m10 : lc == 2 ? m20 : m30;
        m[1] = lc == 0 ? 0.0 : lc == 1 ? i1 : lc == 2 ? m21 : m31;
        m[2] = lc < 2 ? 0.0 : lc == 2 ? i2 : m32;
        m[3] = lc < 3 ? 0.0 : i3;
    }
    // ---- (5) row k of L^-1 H_g over the support, and (L^-1 y)[k]
    double ht[7];
#pragma unroll
    for (int e = 0; e < 7; ++e)
        ht[e] = fma(m[3], s.Hd[3][e], fma(m[2], s.Hd[2][e], fma(m[1], s.Hd[1][e], m[0] * s.Hd[0][e])));
    const double yt = fma(m[3], s.T[3][7], fma(m[2], s.T[2][7], fma(m[1], s.T[1][7], m[0] * s.T[0][7])));
    // ---- (6) W[k][8t + lr], t = 0..4: this lane's A and B operands
    double w[NT];
#pragma unroll
    for (int t = 0; t < NT; ++t) {
        double acc = 0.0;
#pragma unroll
        for (int e = 0; e < 7; ++e)
            if (e < nsup) acc = fma(ht[e], s.Praw[e][8 * t + lr], acc);
        w[t] = acc;
    }
    // column / row 35: no row of P there (A operand 0); as a column it accumulates dx += W' (L^-1 y)
    const double a4 = lr == 3 ? 0.0 : w[NT - 1];
    const double b4 = lr == 3 ? -yt : w[NT - 1];
    // ---- (7) P -= W' W
#pragma unroll
    for (int tr = 0; tr < NT; ++tr) {
        const double an = tr == NT - 1 ? -a4 : -w[tr];
#pragma unroll
        for (int tc = 0; tc < NT; ++tc) dmma(c[tr][tc], an, tc == NT - 1 ? b4 : w[tc]);
    }
    __syncwarp();  // Praw / Hd / T are rewritten by the next block
}

__global__ void __launch_bounds__(WPB * 32, 3) ekf_batched_mma_kernel(const EkfParams kp) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    MmaSlot* slots = reinterpret_cast<MmaSlot*>(smem_raw);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int lr = lane >> 2, lc = lane & 3;
    const uint32_t f = blockIdx.x * WPB + warp;
    if (f >= kp.st.batch) return;  // whole warp; the kernel has no CTA barrier
    MmaSlot& s = slots[warp];

    double2 c[NT][NT];
    {
        const double* Pg = kp.st.P + (size_t)f * NPG * NPG;
#pragma unroll
        for (int tr = 0; tr < NT; ++tr) {
            const int i = 8 * tr + lr;
#pragma unroll
            for (int tc = 0; tc < NT; ++tc) {
                const int j = 8 * tc + 2 * lc;
                const bool in = i < 35;
                c[tr][tc].x = (in && j < 35) ? Pg[to_ref(i) * NPG + to_ref(j)] : 0.0;
                c[tr][tc].y = (in && j + 1 < 35) ? Pg[to_ref(i) * NPG + to_ref(j + 1)] : 0.0;
            }
        }
    }
    for (int j = lane; j < NPG; j += 32) {
        s.X[j] = kp.st.X[(size_t)f * NPG + j];
        s.Z[j] = kp.st.Z[(size_t)f * NPG + j];
        s.Xf[j] = (float)s.X[j];
    }
    if (lane == 0) {
        s.fr.N = kp.st.N[f];
        s.fr.status = kp.st.status[f];
        s.fr.flags = kp.st.flags[f];
        s.fr.wcount = kp.st.wcount[f];
        s.fr.promote = 0;
        s.a00 = kp.st.A[2 * (size_t)f];
        s.a10 = kp.st.A[2 * (size_t)f + 1];
    }
    prefetch_message<32>(s.msg[0], kp.in, kp.st.batch, f, 0, lane);
    cp_async_wait_all();
    __syncwarp();

    for (uint32_t step = 0; step < kp.in.steps; ++step) {
        // ---- cbOdom gate (ekf.cpp:76): nothing happens before the first observation message
        const bool gated = s.fr.flags & FLAG_INIT_Z;
        const MsgBuf& m = s.msg[step & 1];
        if (step + 1 < kp.in.steps) prefetch_message<32>(s.msg[(step + 1) & 1], kp.in, kp.st.batch, f, step + 1, lane);
        if (!gated) {
            // ---- updateZandA, ekf.cpp:139-142
            if (lane == 0) {
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
            for (int l = lane; l < (int)(sizeof(s.fr.owner) / sizeof(int32_t)); l += 32) s.fr.owner[l] = 0;
            __syncwarp();
            front_phase_a<true>(s.fr, kp.fp, s.Xf, s.Z, m.rng, m.brg, 0u, f, lane, 32);
            __syncwarp();
            {
                int32_t* assoc = nullptr;
                if (kp.in.assoc && step + 1 == kp.in.steps)
                    assoc = kp.in.assoc + (kp.in.offsets ? s.fr.obs_base : f * kp.in.obs_per_filter);
                front_phase_b_assoc(s.fr, s.Z, assoc, lane, 32);
            }
            if (lane == 0) {
                front_phase_b_wait(s.fr, kp.fp, s.X, s.Z, f);
                // ekf.cpp:87-91 X = Z on the first processed message: pose rows here, landmark rows below
                const bool initx = s.fr.flags & FLAG_INIT_X;
                motion_jacobian(s.fr.vx, s.fr.az, s.fr.dt, s.Z[2], s.a00, s.a10);  // ekf.cpp:206-212
                // slam(): ekf.cpp:295-296
                double x = initx ? s.Z[0] : s.X[0], y = initx ? s.Z[1] : s.X[1], th = initx ? s.Z[2] : s.X[2];
                motion_model(x, y, th, (float)s.fr.vx, (float)s.fr.az, s.fr.dt);
                s.X[0] = x;
                s.X[1] = y;
                s.X[2] = (double)normalize_angle((float)th, kp.fp.pi);
            }
            __syncwarp();
            const uint32_t N = s.fr.N;
            if (s.fr.flags & FLAG_INIT_X)
                for (uint32_t j = 3 + lane; j < N; j += 32) s.X[j] = s.Z[j];
            // ---- growth of P (ekf.cpp:277): new diagonal entries, the rest of the new rows/cols is 0
            if (s.fr.grew_from) {
                const int g0 = (int)s.fr.grew_from - 3, g1 = (int)N - 3;
#pragma unroll
                for (int t = 0; t < NT - 1; ++t) {
                    const int i = 8 * t + lr;
                    if (i >= g0 && i < g1) {
                        if (2 * lc == lr) c[t][t].x += kp.kp_lm;
                        if (2 * lc + 1 == lr) c[t][t].y += kp.kp_lm;
                    }
                }
            }
            // ---- P = A P A' + Q (ekf.cpp:297); A = I except A(0,0), A(1,0). Pose rows/columns are internal 32, 33.
            {
                const double a00 = s.a00, a10 = s.a10;
#pragma unroll
                for (int tc = 0; tc < NT; ++tc) {  // rows: lanes lr = 0 (row 0) and lr = 1 (row 1) of tile row 4
                    const double x0 = __shfl_sync(0xffffffffu, c[NT - 1][tc].x, lc);
                    const double y0 = __shfl_sync(0xffffffffu, c[NT - 1][tc].y, lc);
                    if (lr == 0) {
                        c[NT - 1][tc].x = a00 * c[NT - 1][tc].x;
                        c[NT - 1][tc].y = a00 * c[NT - 1][tc].y;
                    } else if (lr == 1) {
                        c[NT - 1][tc].x = a10 * x0 + c[NT - 1][tc].x;
                        c[NT - 1][tc].y = a10 * y0 + c[NT - 1][tc].y;
                    }
                }
                if (lc == 0) {  // columns 32 (.x) and 33 (.y) of tile column 4
#pragma unroll
                    for (int tr = 0; tr < NT; ++tr) {
                        const double x0 = c[tr][NT - 1].x;
                        c[tr][NT - 1].x = x0 * a00;
                        c[tr][NT - 1].y = x0 * a10 + c[tr][NT - 1].y;
                    }
                }
                if (lr == 0 && lc == 0) c[NT - 1][NT - 1].x += kp.kq;
                if (lr == 1 && lc == 0) c[NT - 1][NT - 1].y += kp.kq;
                if (lr == 2 && lc == 1) c[NT - 1][NT - 1].x += kp.kq;
            }
            __syncwarp();
            // ---- updateH (ekf.cpp:117-134) and Y = Z - h(X) (ekf.cpp:302-307), one landmark per lane
            {
                const double x = s.X[0], y = s.X[1], th = s.X[2];
                for (uint32_t l = lane; 3 + 2 * l < N; l += 32) {
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
                if (lane == 31) {
                    s.fr.flags &= ~FLAG_INIT_X;
                    s.Y0[0] = s.Z[0] - x;
                    s.Y0[1] = s.Z[1] - y;
                    s.Y0[2] = (double)normalize_angle((float)(s.Z[2] - th), kp.fp.pi);
                }
            }
            __syncwarp();
            // ---- the pose block, then one block per landmark pair
            block_update<0>(c, s, kp, -1, N, lane);
            const int nb = ((int)N - 3 + 3) / 4;
            for (int b = 0; b < nb; ++b) {
                switch (b >> 1) {
                    case 0: block_update<0>(c, s, kp, b, N, lane); break;
                    case 1: block_update<1>(c, s, kp, b, N, lane); break;
                    case 2: block_update<2>(c, s, kp, b, N, lane); break;
                    default: block_update<3>(c, s, kp, b, N, lane); break;
                }
            }
            // ---- X = X + K Y (ekf.cpp:309): the correction sits in column 35 (lanes lc = 1, .y); clear it
            if (lc == 1) {
#pragma unroll
                for (int tr = 0; tr < NT; ++tr) {
                    const int i = 8 * tr + lr;
                    if (i < 35) {
                        const uint32_t g = (uint32_t)to_ref(i);
                        if (g < N) {
                            const double xn = s.X[g] + c[tr][NT - 1].y;
                            s.X[g] = xn;
                            s.Xf[g] = (float)xn;
                        }
                    }
                    c[tr][NT - 1].y = 0.0;
                }
            }
        }
        cp_async_wait_all();  // each lane waits for its own copies of the next message; the barrier publishes them
        __syncwarp();
    }

    // ---- write back
    {
        double* Pg = kp.st.P + (size_t)f * NPG * NPG;
        double acc = 0.0;  // NaN/Inf anywhere makes the sum non-finite
#pragma unroll
        for (int tr = 0; tr < NT; ++tr) {
            const int i = 8 * tr + lr;
#pragma unroll
            for (int tc = 0; tc < NT; ++tc) {
                const int j = 8 * tc + 2 * lc;
                acc += c[tr][tc].x + c[tr][tc].y;
                if (i < 35 && j < 35) Pg[to_ref(i) * NPG + to_ref(j)] = c[tr][tc].x;
                if (i < 35 && j + 1 < 35) Pg[to_ref(i) * NPG + to_ref(j + 1)] = c[tr][tc].y;
            }
        }
        bool bad = !isfinite(acc);
        for (int j = lane; j < NPG; j += 32) {
            kp.st.X[(size_t)f * NPG + j] = s.X[j];
            kp.st.Z[(size_t)f * NPG + j] = s.Z[j];
            bad |= !isfinite(s.X[j]);
        }
        if (bad) atomicOr(&kp.st.status[f], (uint32_t)ASLAM_ST_NONFINITE);
        if (lane == 0) {
            kp.st.N[f] = s.fr.N;
            if (s.fr.status) atomicOr(&kp.st.status[f], s.fr.status);
            kp.st.flags[f] = s.fr.flags;
            kp.st.wcount[f] = s.fr.wcount;
            kp.st.A[2 * (size_t)f] = s.a00;
            kp.st.A[2 * (size_t)f + 1] = s.a10;
        }
    }
}

}  // namespace

aslam_status launch_ekf_batched_mma36(const EkfParams& kp, cudaStream_t stream) {
    const size_t smem = sizeof(MmaSlot) * WPB;
    ASLAM_DYN_SMEM(ekf_batched_mma_kernel, smem);
    const uint32_t grid = (kp.st.batch + WPB - 1) / WPB;
    ekf_batched_mma_kernel<<<grid, WPB * 32, smem, stream>>>(kp);
    ASLAM_CUDA(cudaGetLastError());
    return ASLAM_OK;
}

}  // namespace aslam
