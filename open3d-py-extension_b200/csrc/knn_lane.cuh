// knn_lane.cuh — tiers 2 and 3 of estimate_normals (reference call: PointCloud.py:68-73); tier 1 is knn_select.cuh.
//
// Tier 2 (knn_range_kernel): the queries tier 1 could not certify, compacted in curve order; a warp takes 32 of them and
//   answers one at a time: exact FP64 range query (d2 <= cap, the bound tier 1 left) over the range tables into a shared
//   list, exact (d2, index) ranking, FP64 cumulants; knn_range_finish_kernel (thread per query) turns the cumulants into
//   normals.  List overflow -> tier 3.
// Tier 3 (knn_exact_kernel): warp per query, the exact FP64 sorted-list search of knn.cuh, nearest child first.
//   (Measured and not kept: a middle tier with a 1024-entry (d2, position) list and an 8-pass radix select over the FP64 bit
//   patterns answered 36.6 k of the 49 k overflowing queries of the 10 M scan but left the 12 k with more than 1024 points inside
//   their bound, which are the expensive ones: 1.21 ms against 0.86 ms of the exact search alone.
//   Also measured and not kept: tier 2 batched per warp (ONE walk with the bounding box of the 32 balls, every listed cell
//   group loaded once and tested against all queries of the batch).  The box of 32 curve-consecutive failed queries touches
//   87 cell groups on average where one ball touches ~11, a third of the batches exceed 192: 5.74 ms against 2.19 ms.)
#pragma once
#include "knn_select.cuh"

// ---- tier 2: warp per failed query, exact FP64 range query within cap --------------------------------
// A warp takes up to 32 consecutive entries of the compacted failure list and answers them one at a time.  The bound cap
// is fixed, so the order in which boxes are opened does not matter and the walk runs level by level: the frontier of one
// table is tested K2_NUNR nodes at a time (lane = child box, one 32-byte record each), the surviving children are the next
// frontier, the surviving entries of table 0 a list of cell groups, which are then loaded K2_LUNR at a time (lane = point,
// one 32-byte record each).  Loads of one step do not depend on each other, so a query costs about one memory round trip
// per level and per K2_LUNR cell groups; the depth-first form it replaces (one box, one cell group at a time, the original
// index fetched per hit) paid ~35 dependent round trips per query and was bound by them (ncu r02v: 5.5 warps stalled on
// the long scoreboard per issue, issue slots 51 % busy).  With the bit-pattern ranking below: 2.19 -> 1.91 ms at 10 M.
// Every point with d2 <= cap goes to a shared list; the list is ranked by (d2, original index), the cumulants of the K
// smallest are summed and parked in lane i, and at the end every lane finishes one query.  A frontier, group list or hit
// list that overflows sends the query to tier 3.
#ifndef K2_WARPS
#define K2_WARPS 4
#endif
#ifndef K2_LIST
#define K2_LIST 128
#endif
static_assert(K2_LIST % 32 == 0 && K2_LIST <= 256, "the ranks of the hit list live in K2_LIST / 32 registers per lane");
#ifndef K2_FRONT
#define K2_FRONT 64         // frontier nodes per level
#endif
#ifndef K2_LEAVES
#define K2_LEAVES 128       // cell groups one ball may touch
#endif
#ifndef K2_NUNR
#define K2_NUNR 2
#endif
#ifndef K2_LUNR
#define K2_LUNR 2
#endif
#ifndef K2_MINB
#define K2_MINB 7           // resident blocks per SM the register budget is set for
#endif

struct K2WarpMem {
    static constexpr int LIST = K2_LIST;
    double d[K2_LIST], x[K2_LIST], y[K2_LIST], z[K2_LIST];
    unsigned oidx[K2_LIST];
    unsigned pos[K2_LIST];      // sorted position (neighbour-set record, keep mode)
    int front[2][K2_FRONT];
    int leaf[K2_LEAVES];
};

// lower bound (FP32, every step rounded down, boxes stored rounded outwards) of the squared distance from q to box c
__device__ __forceinline__ float kl_box_lb(const GridDesc &g, int lvl, int c, float qxl, float qxh, float qyl, float qyh, float qzl, float qzh)
{
    const float4 *e = g.boxq + 2 * (g.qoff[lvl] + c);          // one 32-byte record per box: two loads, one address
    const float4 a = e[0];
    const float2 b = *(const float2 *)(e + 1);
    float gx = fmaxf(fmaxf(__fsub_rd(a.x, qxh), __fsub_rd(qxl, a.w)), 0.f);
    float gy = fmaxf(fmaxf(__fsub_rd(a.y, qyh), __fsub_rd(qyl, b.x)), 0.f);
    float gz = fmaxf(fmaxf(__fsub_rd(a.z, qzh), __fsub_rd(qzl, b.y)), 0.f);
    return __fadd_rd(__fadd_rd(__fmul_rd(gx, gx), __fmul_rd(gy, gy)), __fmul_rd(gz, gz));
}

// the range query of one ball: fills sm.d/x/y/z/pos with every point inside, returns their number
// (> K2_LIST: overflow somewhere, the list is not usable)
template <int MODE>
__device__ __forceinline__ int k2_range_query(const GridDesc &g, K2WarpMem &sm, int lane, double qx, double qy, double qz,
                                              double cap, double r2)
{
    const float qxl = __double2float_rd(qx), qxh = __double2float_ru(qx), qyl = __double2float_rd(qy), qyh = __double2float_ru(qy),
                qzl = __double2float_rd(qz), qzh = __double2float_ru(qz), capf = __double2float_ru(cap);
    const unsigned lt = (1u << lane) - 1u;
    int *cur = sm.front[0], *nxt = sm.front[1];
    int np = 1;
    if (lane == 0) cur[0] = 0;
    __syncwarp();
    for (int lvl = g.levels - 1; lvl >= 0; --lvl) {
        int *dst = lvl > 0 ? nxt : sm.leaf;
        const int dcap = lvl > 0 ? K2_FRONT : K2_LEAVES, ccnt = (int)g.cnt[lvl];
        int nn = 0;
        for (int j = 0; j < np; j += K2_NUNR) {
            int c[K2_NUNR];
            float lb[K2_NUNR];
#pragma unroll
            for (int u = 0; u < K2_NUNR; ++u) {
                c[u] = j + u < np ? cur[j + u] * 32 + lane : ccnt;
                lb[u] = c[u] < ccnt ? kl_box_lb(g, lvl, c[u], qxl, qxh, qyl, qyh, qzl, qzh) : KF_HUGE;
            }
#pragma unroll
            for (int u = 0; u < K2_NUNR; ++u) {
                const bool in = lb[u] <= capf;
                const unsigned mk = __ballot_sync(FULL, in);
                const int at = nn + __popc(mk & lt);
                if (in && at < dcap) dst[at] = c[u];
                nn += __popc(mk);
            }
        }
        if (nn > dcap) return K2_LIST + 1;
        __syncwarp();
        int *t = cur; cur = nxt; nxt = t;
        np = nn;
    }
    // np cell groups in sm.leaf
    int cnt = 0;
    const int n = (int)g.n;                                    // n < 2^31: original indices are int32
    for (int s = 0; s < np; s += K2_LUNR) {
        int p[K2_LUNR];
        double4 c4[K2_LUNR];
#pragma unroll
        for (int u = 0; u < K2_LUNR; ++u) {
            p[u] = s + u < np ? sm.leaf[s + u] * 32 + lane : n;
            if (p[u] < n) c4[u] = ld_rec(g.sp4 + p[u]);
        }
#pragma unroll
        for (int u = 0; u < K2_LUNR; ++u) {
            bool hit = false;
            double dd = 0.0;
            if (p[u] < n) {
                dd = dist2_rn(qx, qy, qz, c4[u].x, c4[u].y, c4[u].z);
                hit = dd <= cap && (MODE != 2 || dd < r2);
            }
            const unsigned mk = __ballot_sync(FULL, hit);
            const int at = cnt + __popc(mk & lt);
            if (hit && at < K2_LIST) { sm.d[at] = dd; sm.x[at] = c4[u].x; sm.y[at] = c4[u].y; sm.z[at] = c4[u].z; sm.pos[at] = (unsigned)p[u]; }
            cnt += __popc(mk);
        }
        if (cnt > K2_LIST) return cnt;
    }
    __syncwarp();
    return cnt;
}

template <int MODE>
__global__ void __launch_bounds__(K2_WARPS * 32, K2_MINB)
knn_range_kernel(const GridDesc g, int K, double radius, KnnOut out, const int32_t *__restrict__ list, int m,
                 const float *__restrict__ capbuf, int qpw, double *__restrict__ park)
{
    // park = [11][m]: the nine cumulants, the neighbour count (-1: not answered here) and the k-th d2 of every entry of the
    // failure list; knn_range_finish_kernel turns them into normals, one thread per query (the eigen solver wants more
    // registers than the walk can spare, and one lane of a warp running it 32 times would waste the other 31)
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
    const double r2 = __dmul_rn(radius, radius);
    if (lane < nq) park[(int64_t)9 * m + first + lane] = -1.0;
    __syncwarp();
    for (int i = 0; i < nq; ++i) {
        double cap = shfl_d(mycap, i);
        if (!(cap < 1e300)) continue;            // no bound known: the exact search (tier 3) takes it
        if (MODE == 2 && r2 < cap) cap = r2;
        const int cnt = k2_range_query<MODE>(g, sm, lane, shfl_d(mx, i), shfl_d(my, i), shfl_d(mz, i), cap, r2);
        if (cnt > K2_LIST) continue;             // too many candidates inside the bound: tier 3
        const int64_t qpos = __shfl_sync(FULL, myq, i);
        // the K smallest by (d2, original index) contribute.  Fast form: lt = entries with a smaller d2 (non-negative doubles
        // order like their bit patterns: two integer compares); an entry is in iff lt < K.  Exact unless entries with EQUAL d2
        // straddle the boundary, which shows as more than K members; only then are the original indices fetched and the
        // (d2, index) rank computed.
        Cumul cu;
        cu.clear();
        double kd = 0.0;
        int rk[K2_LIST / 32];                        // rank of entry r * 32 + lane
#pragma unroll
        for (int r = 0; r < K2_LIST / 32; ++r) rk[r] = 0;
        if (cnt > K) {
            int members = 0;
#pragma unroll
            for (int r = 0; r < K2_LIST / 32; ++r) {
                if (r * 32 < cnt) {
                    const int e = r * 32 + lane;
                    const unsigned long long be = e < cnt ? (unsigned long long)__double_as_longlong(sm.d[e]) : ~0ull;
                    int lt = 0;
                    for (int f = 0; f < cnt; ++f) lt += (unsigned long long)__double_as_longlong(sm.d[f]) < be ? 1 : 0;
                    rk[r] = lt;
                    members += __popc(__ballot_sync(FULL, e < cnt && lt < K));
                }
            }
            if (members != K) {
                for (int e = lane; e < cnt; e += 32) sm.oidx[e] = g.sidx[sm.pos[e]];
                __syncwarp();
#pragma unroll
                for (int r = 0; r < K2_LIST / 32; ++r) {
                    if (r * 32 < cnt) {
                        const int e = r * 32 + lane;
                        const bool ve = e < cnt;
                        const double de = ve ? sm.d[e] : 0.0;
                        const unsigned oe = ve ? sm.oidx[e] : 0u;
                        int rank = 0;
                        for (int f = 0; f < cnt; ++f) {
                            double df = sm.d[f];
                            unsigned of = sm.oidx[f];
                            rank += (df < de || (df == de && of < oe)) ? 1 : 0;
                        }
                        rk[r] = rank;
                    }
                }
            }
        }
        int slot = 0;                               // members so far (keep mode: where the next one is recorded)
#pragma unroll
        for (int r = 0; r < K2_LIST / 32; ++r) {
            if (r * 32 < cnt) {
                const int e = r * 32 + lane;
                const bool in = e < cnt && rk[r] < K;
                const unsigned im = __ballot_sync(FULL, in);
                if (in) {
                    cu.add(sm.x[e], sm.y[e], sm.z[e]); kd = fmax(kd, sm.d[e]);
                    if (out.keep) out.nbr[(int64_t)(slot + __popc(im & ((1u << lane) - 1u))) * out.nbr_pitch + qpos] = sm.pos[e];
                }
                slot += __popc(im);
            }
        }
        if (out.keep) {
            for (int j = (cnt < K ? cnt : K) + lane; j < K; j += 32) out.nbr[(int64_t)j * out.nbr_pitch + qpos] = KNN_NONE;
            if (lane == 0) out.tier[qpos] = 2;
        }
#pragma unroll
        for (int dlt = 16; dlt > 0; dlt >>= 1) kd = fmax(kd, shfl_xor_d(kd, dlt));
#pragma unroll
        for (int a = 0; a < 9; ++a) {
            double v = cu.c[a];
#pragma unroll
            for (int dlt = 16; dlt > 0; dlt >>= 1) v += shfl_xor_d(v, dlt);
            if (lane == a) park[(int64_t)a * m + first + i] = v;
        }
        const int tn = __reduce_add_sync(FULL, cu.n);
        if (lane == 9) park[(int64_t)9 * m + first + i] = (double)tn;
        if (lane == 10) park[(int64_t)10 * m + first + i] = kd;
        __syncwarp();
    }
}

template <int MODE>
__global__ void __launch_bounds__(128)
knn_range_finish_kernel(const GridDesc g, KnnOut out, const int32_t *__restrict__ list, int m, float *__restrict__ capbuf,
                        const double *__restrict__ park)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= m) return;
    Cumul mine;
    mine.n = (int)park[(int64_t)9 * m + t];
    if (mine.n < 0) return;
#pragma unroll
    for (int a = 0; a < 9; ++a) mine.c[a] = park[(int64_t)a * m + t];
    const int q = list[t];
    const float kb = fmaxf(__double2float_ru(park[(int64_t)10 * m + t]), 1e-30f);
    const bool stored = out.nx ? kf_store_normal(mine, g.sidx[q], out, KL_GUARD) : true;
    capbuf[q] = stored ? -kb : kb;         // order-sensitive covariance: stays pending for the exact tier, with its bound
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
