// knn_lane.cuh — tiers 2 and 3 of estimate_normals (reference call: PointCloud.py:68-73); tier 1 is knn_select.cuh.
//
// Tier 2 (knn_range_kernel): the queries tier 1 could not certify, compacted in curve order; a warp takes 32 of them and
//   answers one at a time: exact FP64 range query (d2 <= cap, the bound tier 1 left) over the range tables into a shared
//   list, exact (d2, index) ranking, FP64 cumulants, finish.  List overflow -> tier 3.
// Tier 3 (knn_exact_kernel): warp per query, the exact FP64 sorted-list search of knn.cuh, nearest child first.
//   (Measured and not kept: a middle tier with a 1024-entry (d2, position) list and an 8-pass radix select over the FP64 bit
//   patterns answered 36.6 k of the 49 k overflowing queries of the 10 M scan but left the 12 k with more than 1024 points inside
//   their bound, which are the expensive ones: 1.21 ms against 0.86 ms of the exact search alone.)
#pragma once
#include "knn_select.cuh"

// ---- tier 2: warp per failed query, exact FP64 range query within cap --------------------------------
// A warp takes 32 consecutive entries of the compacted failure list.  For each it walks the range tables
// with the fixed bound cap (lane = child box / lane = point, coalesced), appends every point with
// d2 <= cap to a shared list, ranks the list by (d2, original index) and sums the cumulants of the K
// smallest; the totals are parked in lane i, and at the end every lane finishes one query.
#ifndef K2_WARPS
#define K2_WARPS 4
#endif
#ifndef K2_LIST
#define K2_LIST 128
#endif

struct K2WarpMem {
    static constexpr int LIST = K2_LIST;
    double d[K2_LIST], x[K2_LIST], y[K2_LIST], z[K2_LIST];
    unsigned oidx[K2_LIST];
    unsigned pos[K2_LIST];      // sorted position (neighbour-set record, keep mode)
    __device__ __forceinline__ void put(int at, double dd, double xx, double yy, double zz, int64_t p, const GridDesc &g)
    {
        d[at] = dd; x[at] = xx; y[at] = yy; z[at] = zz; oidx[at] = g.sidx[p]; pos[at] = (unsigned)p;
    }
};

// lower bound (FP32, every step rounded down, boxes stored rounded outwards) of the squared distance from q to box c
__device__ __forceinline__ float kl_box_lb(const GridDesc &g, int lvl, int64_t c, float qxl, float qxh, float qyl, float qyh, float qzl, float qzh)
{
    const float *b = g.boxf + g.off[lvl];
    int64_t p = g.pitch[lvl];
    float gx = fmaxf(fmaxf(__fsub_rd(b[c], qxh), __fsub_rd(qxl, b[3 * p + c])), 0.f);
    float gy = fmaxf(fmaxf(__fsub_rd(b[p + c], qyh), __fsub_rd(qyl, b[4 * p + c])), 0.f);
    float gz = fmaxf(fmaxf(__fsub_rd(b[2 * p + c], qzh), __fsub_rd(qzl, b[5 * p + c])), 0.f);
    return __fadd_rd(__fadd_rd(__fmul_rd(gx, gx), __fmul_rd(gy, gy)), __fmul_rd(gz, gz));
}

template <int MODE, class Mem>
struct K2Range {
    const GridDesc &g;
    Mem &sm;
    int lane;
    double qx, qy, qz, cap, r2;
    float fql[3], fqh[3], capf;     // the query rounded down / up and the bound rounded up, for the FP32 box test
    int cnt, K;
    __device__ __forceinline__ K2Range(const GridDesc &g_, Mem &sm_) : g(g_), sm(sm_) {}

    __device__ __forceinline__ void leaf(int64_t leaf_id)
    {
        int64_t p = leaf_id * 32 + lane;
        bool hit = false;
        double x = 0, y = 0, z = 0, d = 0;
        if (p < g.n) {
            x = g.sx[p]; y = g.sy[p]; z = g.sz[p];
            d = dist2_rn(qx, qy, qz, x, y, z);
            hit = d <= cap && (MODE != 2 || d < r2);
        }
        unsigned m = __ballot_sync(FULL, hit);
        int at = cnt + __popc(m & ((1u << lane) - 1u));
        if (hit && at < Mem::LIST) sm.put(at, d, x, y, z, p, g);
        cnt += __popc(m);
    }

    template <int LVL>
    __device__ __forceinline__ void visit(int64_t parent)
    {
        int64_t c = parent * 32 + lane;
        bool ex = c < g.cnt[LVL];
        float lb = ex ? kl_box_lb(g, LVL, c, fql[0], fqh[0], fql[1], fqh[1], fql[2], fqh[2]) : KF_HUGE;
        unsigned m = __ballot_sync(FULL, ex && lb <= capf);
        // fixed bound: every child that intersects the ball is opened, the order does not matter.  (Opening the nearest child
        // first and shrinking the bound to the k-th distance whenever the list fills was measured twice: as a replacement, and
        // as a middle tier for the 49 k overflowing queries of the 10 M scan, 2.2 ms against 1.2 ms of the exact search.)
        while (m && cnt <= Mem::LIST) {
            const int b = __ffs(m) - 1;
            m &= m - 1;
            if constexpr (LVL == 0) leaf(parent * 32 + b);
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

template <int MODE>
__global__ void __launch_bounds__(K2_WARPS * 32)
knn_range_kernel(const GridDesc g, int K, double radius, KnnOut out, const int32_t *__restrict__ list, int m,
                 float *__restrict__ capbuf, int qpw)
{
    // qpw = queries per warp (<= 32): 32 amortises the warp's set-up when there are millions of failed queries; a frame-sized
    // cloud leaves a few thousand, and fewer per warp then means more warps to fill the machine (the host picks it)
    __shared__ K2WarpMem k2_sm[K2_WARPS];
    const int lane = lane_id(), w = threadIdx.x >> 5;
    const int64_t first = ((int64_t)blockIdx.x * K2_WARPS + w) * qpw;
    if (first >= m) return;
    const int nq = (int)(m - first < qpw ? m - first : qpw);
    K2WarpMem &sm = k2_sm[w];
    // lane i owns query i of the batch
    int myq = -1;
    double mx = 0, my = 0, mz = 0, mycap = pos_inf();
    if (lane < nq) {
        myq = list[first + lane];
        mx = g.sx[myq]; my = g.sy[myq]; mz = g.sz[myq];
        float cf = capbuf[myq];
        if (cf < 3.0e38f) mycap = (double)cf;
    }
    K2Range<MODE, K2WarpMem> rq(g, sm);
    rq.lane = lane; rq.K = K;
    rq.r2 = __dmul_rn(radius, radius);
    Cumul mine;
    mine.clear();
    bool done = false;
    double mykth = 0.0;
    for (int i = 0; i < nq; ++i) {
        double cap = shfl_d(mycap, i);
        if (!(cap < 1e300)) continue;            // no bound known: the exact search (tier 3) takes it
        if (MODE == 2 && rq.r2 < cap) cap = rq.r2;
        rq.cap = cap;
        rq.qx = shfl_d(mx, i); rq.qy = shfl_d(my, i); rq.qz = shfl_d(mz, i);
        rq.cnt = 0;
        rq.fql[0] = __double2float_rd(rq.qx); rq.fqh[0] = __double2float_ru(rq.qx);
        rq.fql[1] = __double2float_rd(rq.qy); rq.fqh[1] = __double2float_ru(rq.qy);
        rq.fql[2] = __double2float_rd(rq.qz); rq.fqh[2] = __double2float_ru(rq.qz);
        rq.capf = __double2float_ru(cap);
        rq.walk();
        const int cnt = rq.cnt;
        if (cnt > K2_LIST) continue;             // too many candidates inside the bound: tier 3
        __syncwarp();
        const int64_t qpos = __shfl_sync(FULL, myq, i);
        // rank by (d2, original index); the K smallest contribute
        Cumul cu;
        cu.clear();
        double kd = 0.0;
        for (int e0 = 0; e0 < cnt; e0 += 32) {
            const int e = e0 + lane;
            const bool ve = e < cnt;
            const double de = ve ? sm.d[e] : 0.0;
            const unsigned oe = ve ? sm.oidx[e] : 0u;
            int rank = 0;
            if (cnt > K)
                for (int f = 0; f < cnt; ++f) {
                    double df = sm.d[f];
                    unsigned of = sm.oidx[f];
                    rank += (df < de || (df == de && of < oe)) ? 1 : 0;
                }
            if (ve && rank < K) {
                cu.add(sm.x[e], sm.y[e], sm.z[e]); kd = fmax(kd, de);
                if (out.keep) out.nbr[(int64_t)(cnt > K ? rank : e) * out.nbr_pitch + qpos] = sm.pos[e];     // ranks are distinct
            }
        }
        if (out.keep) {
            for (int j = (cnt < K ? cnt : K) + lane; j < K; j += 32) out.nbr[(int64_t)j * out.nbr_pitch + qpos] = KNN_NONE;
            if (lane == 0) out.tier[qpos] = 2;
        }
#pragma unroll
        for (int dlt = 16; dlt > 0; dlt >>= 1) kd = fmax(kd, shfl_xor_d(kd, dlt));
        if (lane == i) mykth = kd;
#pragma unroll
        for (int a = 0; a < 9; ++a) {
            double v = cu.c[a];
#pragma unroll
            for (int dlt = 16; dlt > 0; dlt >>= 1) v += shfl_xor_d(v, dlt);
            if (lane == i) mine.c[a] = v;
        }
        int tn = __reduce_add_sync(FULL, cu.n);
        if (lane == i) { mine.n = tn; done = true; }
        __syncwarp();
    }
    if (!done) return;
    const float kb = fmaxf(__double2float_ru(mykth), 1e-30f);
    const bool stored = out.nx ? kf_store_normal(mine, g.sidx[myq], out, KL_GUARD) : true;
    capbuf[myq] = stored ? -kb : kb;       // order-sensitive covariance: stays pending for the exact tier, with its bound
}


// ---- tier 3: warp per query, exact FP64 list search -------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(128)
knn_exact_kernel(const GridDesc g, int K, double radius, KnnOut out, const int32_t *__restrict__ list, int m,
                 float *__restrict__ capbuf)
{
    const int w = blockIdx.x * 4 + (threadIdx.x >> 5);
    if (w >= m) return;
    const int q = list[w];
    const float capf = capbuf[q];
    const double cap = capf < 3.0e38f ? (double)capf : pos_inf();
    kf_slow_query<MODE>(g, (int64_t)(q >> 5), q & 31, K, __dmul_rn(radius, radius), cap, out, capbuf + q);
}
