// knn_lane.cuh — estimate_normals, lane-per-query selection (reference call: PointCloud.py:68-73).
//
// Tier 1 (knn_lane_kernel): one warp owns one cell group = 32 Hilbert-consecutive points = 32 queries.
//   STAGE (warp-cooperative, once): the home group, then every point of the home box grown by rho, is
//     written to shared memory as float4 (x, y, z relative to the home-box centre, sorted position).
//   SELECT (every lane answers ITS OWN query; all lanes walk the staged candidates in lockstep, so the
//     candidate is one broadcast LDS.128 and there is no divergence in the two hot loops):
//       pass B  d2 of every candidate -> lane-private 32-bin histogram over [0, U], U = the largest d2
//               to a member of the home group (a valid upper bound of the k-th d2: the group holds >= k
//               points);
//       bin     the bin b* in which the cumulative count crosses k;
//       pass C  candidates below b* are neighbours; those in b* go to a short lane-private list from
//               which the remaining k - c0 are picked; the smallest d2 above the selection is tracked.
//   CERTIFY  with a = k-th and b = (k+1)-th FP32 d2: the set is the exact FP64 (d2, index) set when
//     b - a > 2E (E bounds |d2_fp32 - d2_fp64|) and complete when a + E <= rho^2.  Lanes that cannot
//     certify publish cap = a + E (an upper bound of the true k-th d2) and leave the query to tier 2.
//   FINISH   FP64 cumulants over the selected neighbours, closed-form eigenvector, store.
// Tier 2 (knn_range_kernel): thread per failed query, exact FP64 range query (d2 <= cap) over the range
//   tables into a lane-private list, exact (d2, index) selection, finish.  List overflow -> tier 3.
// Tier 3 (knn_exact_kernel): warp per query, the exact FP64 list search of knn.cuh.
#pragma once
#include "knn_fast.cuh"

#define KL_WARPS 4
#define KL_CAP 768      // staged candidates per warp
#define KL_NB 32        // histogram bins
#define KL_BND 16       // boundary-bin list entries per query

struct KlWarpMem {
    float4 cand[KL_CAP];
    union {
        unsigned hist[KL_NB * 32];                                           // [bin][lane]
        struct { float d[KL_BND * 32]; unsigned short s[KL_BND * 32]; } bnd; // [entry][lane]
    } u;
    unsigned short nb[32 * 32];                                              // [j][lane] staged slots
};

struct KlStage {
    const GridDesc &g;
    KlWarpMem &sm;
    int lane;
    int64_t home;
    double blo[3], bhi[3], cc[3];
    double rho, rho2;
    int n;
    bool overflow;
    unsigned scanned;
    __device__ __forceinline__ KlStage(const GridDesc &g_, KlWarpMem &sm_) : g(g_), sm(sm_) {}

    __device__ __forceinline__ double gap(int lvl, int64_t c) const
    {
        const double *b = g.box + g.off[lvl];
        int64_t p = g.pitch[lvl];
        double gx = fmax(fmax(b[c] - bhi[0], blo[0] - b[3 * p + c]), 0.0);
        double gy = fmax(fmax(b[p + c] - bhi[1], blo[1] - b[4 * p + c]), 0.0);
        double gz = fmax(fmax(b[2 * p + c] - bhi[2], blo[2] - b[5 * p + c]), 0.0);
        return (gx * gx + gy * gy) + gz * gz;
    }

    __device__ __forceinline__ void scan_group(int64_t leaf)
    {
        int64_t p = leaf * 32 + lane;
        bool in = false;
        double lx = 0, ly = 0, lz = 0;
        if (p < g.n) {
            double x = g.sx[p], y = g.sy[p], z = g.sz[p];
            lx = x - cc[0]; ly = y - cc[1]; lz = z - cc[2];
            in = x >= blo[0] - rho && x <= bhi[0] + rho && y >= blo[1] - rho && y <= bhi[1] + rho &&
                 z >= blo[2] - rho && z <= bhi[2] + rho;
        }
        unsigned m = __ballot_sync(FULL, in);
        int at = n + __popc(m & ((1u << lane) - 1u));
        if (in && at < KL_CAP) sm.cand[at] = make_float4((float)lx, (float)ly, (float)lz, __int_as_float((int)p));
        n += __popc(m);
        if (n > KL_CAP) overflow = true;
        scanned++;
    }

    template <int LVL>
    __device__ __forceinline__ void visit(int64_t parent)
    {
        int64_t c = parent * 32 + lane;
        bool ex = c < g.cnt[LVL];
        double gp = ex ? gap(LVL, c) : pos_inf();
        unsigned m = __ballot_sync(FULL, ex && gp <= rho2);
        while (m && !overflow) {
            int b = __ffs(m) - 1;
            m &= m - 1;
            if constexpr (LVL == 0) { if (parent * 32 + b != home) scan_group(parent * 32 + b); }
            else visit<LVL - 1>(parent * 32 + b);
        }
    }

    __device__ __forceinline__ void walk()
    {
        switch (g.levels - 1) {
        case 0: visit<0>(0); break;
        case 1: visit<1>(0); break;
        case 2: visit<2>(0); break;
        case 3: visit<3>(0); break;
        case 4: visit<4>(0); break;
        default: visit<5>(0); break;
        }
    }
};

__device__ __forceinline__ float kl_d2(const float4 c, float qx, float qy, float qz)
{
    float dx = c.x - qx, dy = c.y - qy, dz = c.z - qz;
    return fmaf(dz, dz, fmaf(dy, dy, dx * dx));
}

// stats slots (out.stats, O3DPP_STATS=1): 16 warps, 17 candidates, 18 groups scanned, 19 fail-gap,
// 20 fail-rho, 21 fail-boundary-overflow, 22 fail-count, 23 fail-warp (unstaged / short group), 24 restages
template <int MODE>
__global__ void __launch_bounds__(KL_WARPS * 32)
knn_lane_kernel(const GridDesc g, int K, double radius, float rho_scale, KnnOut out, float *__restrict__ capbuf)
{
    extern __shared__ __align__(16) unsigned char kl_smem[];
    const int lane = lane_id(), w = threadIdx.x >> 5;
    const int64_t leaf_id = (int64_t)blockIdx.x * KL_WARPS + w;
    if (leaf_id >= g.cnt[0]) return;
    KlWarpMem &sm = ((KlWarpMem *)kl_smem)[w];
    const int64_t base = leaf_id * 32;
    const int nq = (int)(g.n - base < 32 ? g.n - base : 32);
    const int64_t mypos = base + lane;
    const bool valid = lane < nq;
    const double r2d = __dmul_rn(radius, radius);
    double px = 0, py = 0, pz = 0;
    if (valid) { px = g.sx[mypos]; py = g.sy[mypos]; pz = g.sz[mypos]; }

    // ---- stage ---------------------------------------------------------------------------------
    KlStage sg(g, sm);
    sg.lane = lane; sg.scanned = 0; sg.home = leaf_id;
    {
        const double *b0 = g.box + g.off[0];
        int64_t p0 = g.pitch[0];
        sg.blo[0] = b0[leaf_id]; sg.blo[1] = b0[p0 + leaf_id]; sg.blo[2] = b0[2 * p0 + leaf_id];
        sg.bhi[0] = b0[3 * p0 + leaf_id]; sg.bhi[1] = b0[4 * p0 + leaf_id]; sg.bhi[2] = b0[5 * p0 + leaf_id];
    }
    sg.cc[0] = 0.5 * (sg.blo[0] + sg.bhi[0]); sg.cc[1] = 0.5 * (sg.blo[1] + sg.bhi[1]); sg.cc[2] = 0.5 * (sg.blo[2] + sg.bhi[2]);
    const double ex = sg.bhi[0] - sg.blo[0], ey = sg.bhi[1] - sg.blo[1], ez = sg.bhi[2] - sg.blo[2];
    double rho = sqrt((ex * ex + ey * ey) + ez * ez) * (double)rho_scale;
    if (MODE == 2 && radius * 1.001 < rho) rho = radius * 1.001;
    // own query in local FP32; the home group occupies slots 0..31 (bit-identical to the query copies)
    const float qx = valid ? (float)(px - sg.cc[0]) : KF_BIG, qy = valid ? (float)(py - sg.cc[1]) : KF_BIG,
                qz = valid ? (float)(pz - sg.cc[2]) : KF_BIG;
    sm.cand[lane] = make_float4(qx, qy, qz, __int_as_float(valid ? (int)mypos : -1));
    bool staged = true;
    unsigned restage = 0;
    for (int attempt = 0;; ++attempt) {
        sg.n = 32; sg.overflow = false; sg.rho = rho; sg.rho2 = rho * rho * (1.0 + 1e-12);
        sg.walk();
        if (!sg.overflow) break;
        restage++;
        if (attempt >= 8) { staged = false; sg.n = 32; break; }
        rho *= 0.7;
    }
    const int ncand = sg.n;
    const int np = (ncand + 3) & ~3;
    if (lane < np - ncand) sm.cand[ncand + lane] = make_float4(KF_BIG, KF_BIG, KF_BIG, __int_as_float(-1));
    __syncwarp();

    // error model of the staged FP32 coordinates: |coordinate| <= half extent + rho
    const float Lf = (float)((0.5 * fmax(fmax(ex, ey), ez) + rho) * 1.001);
    const float eA = 16.0f * KF_U * Lf, eC = 2.3e-13f * Lf * Lf;
    const float rho1f = staged ? __double2float_rd(rho * rho * (1.0 - 1e-6)) : -1.0f;
    float tcap = KF_TCAP;
    if (MODE == 2) {
        float r2f = __double2float_ru(r2d);
        tcap = r2f * (1.0f + 1e-6f) + 2.0f * kf_err(r2f, eA, eC);
        if (!(tcap < KF_TCAP)) staged = false;
    }

    // ---- select: lane = query --------------------------------------------------------------------
    const float4 *__restrict__ cd = sm.cand;
    float U = 0.f;
#pragma unroll 8
    for (int s = 0; s < 32; ++s) {
        float d = kl_d2(cd[s], qx, qy, qz);
        if (s < nq) U = fmaxf(U, d);
    }
    bool ok = false;
    float capf = __int_as_float(0x7f800000);
    int cn = 0;
    unsigned why = 0;     // 1 gap, 2 rho, 3 boundary overflow, 4 count, 5 warp
    if (nq >= K && staged) {
        float T = U;
        bool capped = false;
        if (MODE == 2 && tcap < T) { T = tcap; capped = true; }
        const bool dead = !(T > 1e-36f);
        const float scale = dead ? 0.f : (float)KL_NB / (T * 1.0001f);
        unsigned *hist = sm.u.hist + lane;
#pragma unroll
        for (int b = 0; b < KL_NB; ++b) hist[b * 32] = 0u;
        // pass B: histogram of d2 over [0, T]
#pragma unroll 4
        for (int s = 0; s < np; ++s) {
            float d = kl_d2(cd[s], qx, qy, qz);
            int b = __float2int_rz(d * scale);
            if (b < KL_NB) hist[b * 32]++;
        }
        int c0 = 0, bstar = -1, cb = 0, acc = 0;
#pragma unroll
        for (int b = 0; b < KL_NB; ++b) {
            int h = (int)hist[b * 32];
            if (bstar < 0 && acc + h >= K) { bstar = b; c0 = acc; cb = h; }
            acc += h;
        }
        const bool takeall = bstar < 0 && MODE == 2 && capped;   // fewer than K inside the radius
        const bool counted = bstar >= 0;
        if (takeall) bstar = KL_NB;
        __syncwarp();     // the boundary lists alias the histograms of other lanes
        // pass C: collect
        int j = 0, mb = 0;
        float bmin = KF_HUGE;
        float *bd = sm.u.bnd.d + lane;
        unsigned short *bs = sm.u.bnd.s + lane, *nbl = sm.nb + lane;
#pragma unroll 4
        for (int s = 0; s < np; ++s) {
            float d = kl_d2(cd[s], qx, qy, qz);
            int b = __float2int_rz(d * scale);
            if (b < bstar) { if (j < 32) nbl[j * 32] = (unsigned short)s; j++; }
            else if (b == bstar) { if (mb < KL_BND) { bd[mb * 32] = d; bs[mb * 32] = (unsigned short)s; } mb++; }
            else bmin = fminf(bmin, d);
        }
        if (dead) { why = 4; capf = U + kf_err(U, eA, eC) * 1.0001f; }
        else if (takeall) {
            ok = tcap + kf_err(tcap, eA, eC) <= rho1f && j <= 32;
            capf = tcap;
            if (!ok) why = 2;
        } else if (!counted) { why = 4; capf = U + kf_err(U, eA, eC) * 1.0001f; }
        else if (cb > KL_BND) {
            why = 3;
            float edge = (float)(bstar + 1) / scale * 1.001f;
            capf = edge + kf_err(edge, eA, eC) * 1.0001f;
        } else {
            const int need = K - c0;
            float a = 0.f;
            for (int r = 0; r < need; ++r) {
                float best = KF_HUGE;
                int be = 0;
                for (int e = 0; e < cb; ++e) { float v = bd[e * 32]; if (v < best) { best = v; be = e; } }
                a = best;
                if (j < 32) nbl[j * 32] = bs[be * 32];
                j++;
                bd[be * 32] = KF_HUGE;
            }
            float rest = bmin;
            for (int e = 0; e < cb; ++e) rest = fminf(rest, bd[e * 32]);
            const float ea = kf_err(a, eA, eC);
            const bool c1 = rest - a > 2.0f * kf_err(rest, eA, eC), c2 = a + ea <= rho1f;
            ok = c1 && c2;
            capf = a + ea * 1.0001f;
            if (!c1) why = 1; else if (!c2) why = 2;
        }
        cn = j;
    } else {
        why = 5;
        if (nq >= K) capf = U + kf_err(U, eA, eC) * 1.0001f;
    }
    if (valid) capbuf[mypos] = ok ? -1.0f : capf;
    if (out.stats) {
        unsigned m1 = __ballot_sync(FULL, valid && why == 1), m2 = __ballot_sync(FULL, valid && why == 2), m3 = __ballot_sync(FULL, valid && why == 3),
                 m4 = __ballot_sync(FULL, valid && why == 4), m5 = __ballot_sync(FULL, valid && why == 5);
        if (lane == 0) {
            atomicAdd(out.stats + 16, 1ull);
            atomicAdd(out.stats + 17, (unsigned long long)ncand);
            atomicAdd(out.stats + 18, (unsigned long long)sg.scanned);
            atomicAdd(out.stats + 19, (unsigned long long)__popc(m1));
            atomicAdd(out.stats + 20, (unsigned long long)__popc(m2));
            atomicAdd(out.stats + 21, (unsigned long long)__popc(m3));
            atomicAdd(out.stats + 22, (unsigned long long)__popc(m4));
            atomicAdd(out.stats + 23, (unsigned long long)__popc(m5));
            atomicAdd(out.stats + 24, (unsigned long long)restage);
        }
    }
    // ---- finish ------------------------------------------------------------------------------------
    if (!valid || !ok || !out.nx) return;
    Cumul cu;
    cu.clear();
    for (int j = 0; j < cn; ++j) {
        int p = __float_as_int(cd[sm.nb[j * 32 + lane]].w);
        double x = g.sx[p], y = g.sy[p], z = g.sz[p];
        if (MODE == 2 && !(dist2_rn(px, py, pz, x, y, z) < r2d)) continue;
        cu.add(x, y, z);
    }
    kf_store_normal(cu, g.sidx[mypos], out);
}

// ---- tier 2: thread per failed query, exact FP64 range query within cap -----------------------------
#define K2_THREADS 64
#define K2_LIST 64

__device__ __forceinline__ double kl_box_lb(const GridDesc &g, int lvl, int64_t c, double qx, double qy, double qz)
{
    const double *b = g.box + g.off[lvl];
    int64_t p = g.pitch[lvl];
    double gx = fmax(fmax(__dsub_rn(b[c], qx), __dsub_rn(qx, b[3 * p + c])), 0.0);
    double gy = fmax(fmax(__dsub_rn(b[p + c], qy), __dsub_rn(qy, b[4 * p + c])), 0.0);
    double gz = fmax(fmax(__dsub_rn(b[2 * p + c], qz), __dsub_rn(qz, b[5 * p + c])), 0.0);
    return __dadd_rn(__dadd_rn(__dmul_rn(gx, gx), __dmul_rn(gy, gy)), __dmul_rn(gz, gz));
}

template <int MODE>
__global__ void __launch_bounds__(K2_THREADS)
knn_range_kernel(const GridDesc g, int K, double radius, KnnOut out, const int32_t *__restrict__ list, int m,
                 float *__restrict__ capbuf)
{
    extern __shared__ __align__(16) unsigned char k2_smem[];
    double *ld = (double *)k2_smem + threadIdx.x;                              // [entry][thread]
    int *lp = (int *)((double *)k2_smem + K2_LIST * K2_THREADS) + threadIdx.x;
    const int t = blockIdx.x * K2_THREADS + threadIdx.x;
    if (t >= m) return;
    const int q = list[t];
    const float capf = capbuf[q];
    if (!(capf < 3.0e38f)) return;          // no bound known: the exact search (tier 3) takes it
    const double qx = g.sx[q], qy = g.sy[q], qz = g.sz[q];
    const double r2 = __dmul_rn(radius, radius);
    double cap = (double)capf;
    if (MODE == 2 && r2 < cap) cap = r2;
    int cnt = 0;
    int lvl = g.levels - 1;
    int64_t node[KNN_MAX_LEVELS];
    int ch[KNN_MAX_LEVELS];
    node[lvl] = 0; ch[lvl] = 0;
    while (lvl < g.levels) {
        const int k = ch[lvl];
        const int64_t c = node[lvl] * 32 + k;
        if (k >= 32 || c >= g.cnt[lvl]) { ++lvl; continue; }
        ch[lvl] = k + 1;
        if (kl_box_lb(g, lvl, c, qx, qy, qz) > cap) continue;
        if (lvl > 0) { --lvl; node[lvl] = c; ch[lvl] = 0; continue; }
        const int64_t p0 = c * 32;
        const int e = (int)(g.n - p0 < 32 ? g.n - p0 : 32);
        for (int j = 0; j < e; ++j) {
            double d = dist2_rn(qx, qy, qz, g.sx[p0 + j], g.sy[p0 + j], g.sz[p0 + j]);
            if (d <= cap && (MODE != 2 || d < r2)) {
                if (cnt < K2_LIST) { ld[cnt * K2_THREADS] = d; lp[cnt * K2_THREADS] = (int)(p0 + j); }
                cnt++;
            }
        }
        if (cnt > K2_LIST) break;
    }
    if (cnt > K2_LIST) return;              // too many candidates inside the bound: tier 3
    // keep the K smallest by (d2, original index): drop the largest until K remain
    int mc = cnt;
    while (mc > K) {
        double bd = ld[0];
        int be = 0, bp = lp[0];
        for (int e = 1; e < mc; ++e) {
            double d = ld[e * K2_THREADS];
            int p = lp[e * K2_THREADS];
            if (d > bd || (d == bd && g.sidx[p] > g.sidx[bp])) { bd = d; be = e; bp = p; }
        }
        --mc;
        ld[be * K2_THREADS] = ld[mc * K2_THREADS];
        lp[be * K2_THREADS] = lp[mc * K2_THREADS];
    }
    capbuf[q] = -1.0f;
    if (!out.nx) return;
    Cumul cu;
    cu.clear();
    for (int e = 0; e < mc; ++e) {
        int p = lp[e * K2_THREADS];
        cu.add(g.sx[p], g.sy[p], g.sz[p]);
    }
    kf_store_normal(cu, g.sidx[q], out);
}

// ---- tier 3: warp per query, exact FP64 list search -------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(128)
knn_exact_kernel(const GridDesc g, int K, double radius, KnnOut out, const int32_t *__restrict__ list, int m,
                 const float *__restrict__ capbuf)
{
    const int w = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (w >= m) return;
    const int q = list[w];
    const float capf = capbuf[q];
    const double cap = capf < 3.0e38f ? (double)capf : pos_inf();
    kf_slow_query<MODE>(g, (int64_t)(q >> 5), q & 31, K, __dmul_rn(radius, radius), cap, out);
}
