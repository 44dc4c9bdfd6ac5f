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
//   OUTPUT   the K selected sorted positions of every certified lane go to global memory ([j][query position], one
//     coalesced 128-byte store per j); knn_finish_kernel (knn_finish.cuh), a separate full-occupancy launch, turns
//     them into FP64 cumulants and the closed-form eigenvector.
// Tier 2 (knn_range_kernel): thread per failed query, exact FP64 range query (d2 <= cap) over the range
//   tables into a lane-private list, exact (d2, index) selection, finish.  List overflow -> tier 3.
// Tier 3 (knn_exact_kernel): warp per query, the exact FP64 list search of knn.cuh.
#pragma once
#include "knn_finish.cuh"

#ifndef KL_WARPS
#define KL_WARPS 1
#endif
#ifndef KL_CAP
#define KL_CAP 752      // staged candidates per warp (752 + 2 KL_UNR of padding + KL_BND 22 + KL_LIST 56: eleven blocks per SM)
#endif
#ifndef KL_NB
#define KL_NB 32        // histogram bins
#endif
#ifndef KL_GMAX
#define KL_GMAX 512     // cell groups one part may scan
#endif
#ifndef KL_GSPLIT
#define KL_GSPLIT 192   // more groups than this to scan: split the part without trying
#endif
#ifndef KL_UNR
#define KL_UNR 8       // candidates per iteration of the two select passes (10M, after the code-size cuts: 2 -> 11.52 ms, 4 -> 10.66,
                       // 8 -> 10.27, 16 -> 10.60)
#endif
#ifndef KL_SB
#define KL_SB 4         // cell groups loaded per staging batch (3 * KL_SB loads in flight per lane)
#endif
#ifndef KL_PIPE
#define KL_PIPE 0       // select passes: load the next candidates before processing the current ones
#endif
#ifndef KL_ATOM
#define KL_ATOM 1       // histogram update: 1 = shared atomic (fire and forget), 0 = load/add/store
#endif
#ifndef KL_MINB
#define KL_MINB 12
#endif
#ifndef KL_WIN
#define KL_WIN 2        // unstageable parts: cell groups taken before and after the home group along the curve
#endif
#ifndef KL_BND
#define KL_BND 22       // boundary-bin list entries per query
#endif

#ifndef KL_BFS
#define KL_BFS 1        // phase 1 of staging: 1 = level by level (the box loads of a whole level in flight), 0 = depth first
#endif
#ifndef KL_BFB
#define KL_BFB 2        // level-by-level walk: frontier nodes whose child boxes are loaded together, unconditionally through a
                        // clamped index (10M: 1 -> 10.09 ms, 2 -> 9.90, 4 -> 10.15; with a branch around each node's loads 1 -> 10.74,
                        // 2 -> 10.76, 4 -> 11.37, 8 -> 11.44; depth first 10.96)
#endif

#ifndef KL_LIST
#define KL_LIST 56      // survivors of the threshold pass per query (16-bit staged slots): < K + KL_BND
#endif

static_assert(32 - 1 + KL_BND < KL_LIST, "survivor list: K - 1 below the boundary bin + KL_BND inside it (K <= 32 on this path)");

struct KlWarpMem {
    // candidates, two to a 32-byte record {x0 x1 y0 y1 | z0 z1 p0 p1} so that the select passes run on packed
    // FP32 pairs (FFMA2); p = sorted position bits
    float4 cand[KL_CAP + 2 * KL_UNR];
    union {
        unsigned hist[(KL_NB + 1) * 32];                                        // [bin][lane]
        struct { float d[KL_BND * 32]; unsigned short s[KL_BND * 32]; } bnd; // [entry][lane]
    } u;
    union {
        unsigned short nb[KL_LIST * 32];                                     // [j][lane] staged slots: survivors, then the K neighbours
        unsigned gl[KL_GMAX];                                                // cell groups to scan (stage only)
    } v;
};


// candidate s lives in half (s & 1) of record s >> 1
__device__ __forceinline__ void kl_put(float4 *cand, int s, float x, float y, float z, int pos)
{
    float *r = (float *)(cand + (s & ~1)) + (s & 1);
    r[0] = x; r[2] = y; r[4] = z; r[6] = __int_as_float(pos);
}
__device__ __forceinline__ int kl_pos(const float4 *cand, int s)
{
    return __float_as_int(((const float *)(cand + (s & ~1)))[6 + (s & 1)]);
}
// d2 of the two candidates of record {a, b} to the query (nq* = the negated query coordinate, twice)
__device__ __forceinline__ float2 kl_d2pair(const float4 a, const float4 b, float2 nqx, float2 nqy, float2 nqz)
{
    float2 dx = __fadd2_rn(make_float2(a.x, a.y), nqx), dy = __fadd2_rn(make_float2(a.z, a.w), nqy),
           dz = __fadd2_rn(make_float2(b.x, b.y), nqz);
    return __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
}

// d2 of staged slot s, the same three roundings as the packed form
__device__ __forceinline__ float kl_d2slot(const float4 *cand, int s, float qx, float qy, float qz)
{
    const float *r = (const float *)(cand + (s & ~1)) + (s & 1);
    float dx = __fadd_rn(r[0], -qx), dy = __fadd_rn(r[2], -qy), dz = __fadd_rn(r[4], -qz);
    return __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
}

// stores through 32-bit shared addresses: a running generic pointer costs a 64-bit add per step
__device__ __forceinline__ void kl_sts16(unsigned a, unsigned short v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"(v) : "memory"); }
__device__ __forceinline__ void kl_sts32(unsigned a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }

struct KlStage {
    const GridDesc &g;
    KlWarpMem &sm;
    int lane;
    int64_t home;
    double blo[3], bhi[3], cc[3];
    double rho, rho2;
    float flo[3], fhi[3], rho2f;     // the part's box and rho^2 in FP32, rounded outwards / up
    int n, ngroups, gsplit;
    unsigned scanned;
    __device__ __forceinline__ KlStage(const GridDesc &g_, KlWarpMem &sm_) : g(g_), sm(sm_) {}

    // lower bound of the squared box-to-box gap in FP32: the stored boxes are rounded outwards, the part's box
    // (flo, fhi) too, and every operation rounds down, so a group that matters is never culled
    __device__ __forceinline__ float gap(int lvl, int64_t c) const
    {
        const float *b = g.boxf + g.off[lvl];
        int64_t p = g.pitch[lvl];
        float gx = fmaxf(fmaxf(__fsub_rd(b[c], fhi[0]), __fsub_rd(flo[0], b[3 * p + c])), 0.f);
        float gy = fmaxf(fmaxf(__fsub_rd(b[p + c], fhi[1]), __fsub_rd(flo[1], b[4 * p + c])), 0.f);
        float gz = fmaxf(fmaxf(__fsub_rd(b[2 * p + c], fhi[2]), __fsub_rd(flo[2], b[5 * p + c])), 0.f);
        return __fadd_rd(__fadd_rd(__fmul_rd(gx, gx), __fmul_rd(gy, gy)), __fmul_rd(gz, gz));
    }

    // phase 1: the cell groups whose box comes within rho of the part's box (the home group is staged already)
    template <int LVL>
    __device__ __forceinline__ void collect(int64_t parent)
    {
        int64_t c = parent * 32 + lane;
        bool ex = c < g.cnt[LVL];
        float gp = ex ? gap(LVL, c) : KF_HUGE;
        if constexpr (LVL == 0) {
            bool take = ex && gp <= rho2f && c != home;
            unsigned m = __ballot_sync(FULL, take);
            int at = ngroups + __popc(m & ((1u << lane) - 1u));
            if (take && at < KL_GMAX) sm.v.gl[at] = (unsigned)c;
            ngroups += __popc(m);
        } else {
            unsigned m = __ballot_sync(FULL, ex && gp <= rho2f);
            while (m && ngroups <= gsplit) {
                int b = __ffs(m) - 1;
                m &= m - 1;
                collect<LVL - 1>(parent * 32 + b);
            }
        }
    }

    // phase 1, level by level: the frontier (the nodes of one level whose box comes within rho) ping-pongs between the
    // group list and the histogram area (unused while staging); the child boxes of KL_BFB frontier nodes are loaded
    // together, so a level costs one memory latency instead of one per visited node.  Children are appended in node
    // order, which is the depth-first order: the group list is the same.  false = a frontier overflowed (far more
    // groups than gsplit below it): the attempt is abandoned like one with too many groups.
    __device__ __forceinline__ bool collect_levels()
    {
        const int top = g.levels - 1;
        const unsigned lt = (1u << lane) - 1u;
        if (lane == 0) (((top + 1) & 1) ? sm.u.hist : sm.v.gl)[0] = 0u;
        int ncur = 1;
        for (int lvl = top; lvl >= 0; --lvl) {
            __syncwarp();
            const unsigned *cur = ((lvl + 1) & 1) ? sm.u.hist : sm.v.gl;     // level 0 writes the group list
            unsigned *nxt = (lvl & 1) ? sm.u.hist : sm.v.gl;
            const int64_t cnt = g.cnt[lvl];
            int nout = 0;
            const int lim = lvl ? KL_GMAX - 32 * KL_BFB : gsplit;
            int i = 0;
#if KL_BFB > 1
            for (; i + KL_BFB <= ncur && nout <= lim; i += KL_BFB) {       // full batches: 6 * KL_BFB box loads in flight
                int64_t c[KL_BFB];
                float gp[KL_BFB];
#pragma unroll
                for (int k = 0; k < KL_BFB; ++k) {          // unconditional loads (clamped index): all of them in flight together
                    c[k] = (int64_t)cur[i + k] * 32 + lane;
                    gp[k] = gap(lvl, c[k] < cnt ? c[k] : cnt - 1);
                }
#pragma unroll
                for (int k = 0; k < KL_BFB; ++k) if (c[k] >= cnt) gp[k] = KF_HUGE;
#pragma unroll
                for (int k = 0; k < KL_BFB; ++k) {
                    const bool take = gp[k] <= rho2f && (lvl > 0 || c[k] != home);
                    const unsigned m = __ballot_sync(FULL, take);
                    const int at = nout + __popc(m & lt);
                    if (take && at < KL_GMAX) nxt[at] = (unsigned)c[k];
                    nout += __popc(m);
                }
            }
#endif
#pragma unroll 1
            for (; i < ncur && nout <= lim; ++i) {                            // the rest (and the narrow top levels) one by one
                const int64_t c = (int64_t)cur[i] * 32 + lane;
                const float gp = gap(lvl, c < cnt ? c : cnt - 1);
                const bool take = c < cnt && gp <= rho2f && (lvl > 0 || c != home);
                const unsigned m = __ballot_sync(FULL, take);
                const int at = nout + __popc(m & lt);
                if (take && at < KL_GMAX) nxt[at] = (unsigned)c;
                nout += __popc(m);
            }
            if (lvl && nout > KL_GMAX - 32 * KL_BFB) { ngroups = gsplit + 1; return false; }
            ncur = nout;
        }
        ngroups = ncur;
        return true;
    }

    __device__ __forceinline__ void append(int64_t p, double x, double y, double z)
    {
        bool in = p < g.n && x >= blo[0] - rho && x <= bhi[0] + rho && y >= blo[1] - rho && y <= bhi[1] + rho &&
                  z >= blo[2] - rho && z <= bhi[2] + rho;
        unsigned m = __ballot_sync(FULL, in);
        int at = n + __popc(m & ((1u << lane) - 1u));
        if (in && at < KL_CAP) kl_put(sm.cand, at, (float)(x - cc[0]), (float)(y - cc[1]), (float)(z - cc[2]), (int)p);
        n += __popc(m);
    }

    // phase 2: scan the collected groups, four at a time so that twelve loads are in flight per lane.
    // false = the staging area overflowed or there are too many groups (the caller splits the part).
    __device__ __forceinline__ bool stage()
    {
        ngroups = 0;
#pragma unroll
        for (int a = 0; a < 3; ++a) { flo[a] = __double2float_rd(blo[a]); fhi[a] = __double2float_ru(bhi[a]); }
        rho2f = __double2float_ru(rho2);
#if KL_BFS
        if (!collect_levels()) return false;
#else
        switch (g.levels - 1) {
        case 0: collect<0>(0); break;
        case 1: collect<1>(0); break;
        case 2: collect<2>(0); break;
        case 3: collect<3>(0); break;
        case 4: collect<4>(0); break;
        default: collect<5>(0); break;
        }
#endif
        if (ngroups > gsplit) return false;
        __syncwarp();
        n = 32;
        const int64_t last = g.n - 1;
        for (int i = 0; i < ngroups; i += KL_SB) {
            int64_t p[KL_SB];
            double x[KL_SB], y[KL_SB], z[KL_SB];
#pragma unroll
            for (int k = 0; k < KL_SB; ++k) {
                p[k] = i + k < ngroups ? (int64_t)sm.v.gl[i + k] * 32 + lane : g.n;
                int64_t q = p[k] < last ? p[k] : last;
                x[k] = g.sx[q]; y[k] = g.sy[q]; z[k] = g.sz[q];
            }
#pragma unroll
            for (int k = 0; k < KL_SB; ++k) append(p[k], x[k], y[k], z[k]);
            scanned += KL_SB;
            if (n > KL_CAP) return false;
        }
        return true;
    }
};


// quarter-octave bin of a non-negative float relative to `base` (the bits of d2 >> 21 keep the exponent and
// two mantissa bits): monotone in d2, three integer instructions
__device__ __forceinline__ int kl_bin(float d, int base)
{
    int b = (__float_as_int(d) >> 21) - base;
    return min(max(b, 0), KL_NB);          // KL_NB = beyond the range (its counter is never read)
}

// what one lane learnt about its query
struct KlAnswer {
    bool ok;          // certified: sm.v.nb[.][lane] holds cn staged slots
    int cn;
    float capf;       // upper bound of the true k-th d2 when not certified (+inf: none)
    unsigned why;     // 1 gap, 2 rho, 3 boundary overflow, 4 count, 5 not staged
};

// SELECT + CERTIFY of every lane's own query against the np staged candidates (np % KL_UNR == 0).
// U = an upper bound of the k-th d2 of the staged set (the histogram range; the hybrid radius cap if smaller).
//   A  lane-private quarter-octave histogram of d2 over all candidates; b* = the bin that holds the k-th d2
//   C  second pass: d2 below the upper edge of b* -> the slot joins the survivor list (16 bits each; c0 + cb <
//      K + KL_BND entries), otherwise the smallest rejected d2 is tracked: one compare, one store per candidate
//   F  the survivors are looked at again: bins below b* are neighbours (compacted in place), those of b* go to the
//      boundary list and are ranked exactly; certification of the k-th against the (k+1)-th d2 and rho.
template <int MODE>
__device__ __forceinline__ KlAnswer kl_select(KlWarpMem &sm, int lane, int np, int K, float qx, float qy, float qz, float U,
                                              float tcap, float rho1f, float eA, float eC)
{
    KlAnswer A;
    A.ok = false; A.cn = 0; A.why = 0;
    const float4 *cd = sm.cand;
    float T = U;
    bool capped = false;
    if (MODE == 2 && tcap < T) { T = tcap; capped = true; }
    const bool dead = !(T > 1e-36f);
    const int base = (__float_as_int(dead ? 1.0f : T) >> 21) - (KL_NB - 1);
    unsigned *hist = sm.u.hist + lane;
    const float2 nqx = make_float2(-qx, -qx), nqy = make_float2(-qy, -qy), nqz = make_float2(-qz, -qz);
    unsigned short *const list = sm.v.nb + lane;
#pragma unroll
    for (int b = 0; b < KL_NB; ++b) hist[b * 32] = 0u;
    // pass A (candidates are broadcast reads; unconditional shared atomics, bin KL_NB = beyond the range)
    for (int s = 0; s < np; s += KL_UNR) {
        float4 cc[KL_UNR];
        int bb[KL_UNR];
#pragma unroll
        for (int u = 0; u < KL_UNR; ++u) cc[u] = cd[s + u];
#pragma unroll
        for (int u = 0; u < KL_UNR; u += 2) {
            float2 d = kl_d2pair(cc[u], cc[u + 1], nqx, nqy, nqz);
            bb[u] = kl_bin(d.x, base); bb[u + 1] = kl_bin(d.y, base);
        }
#pragma unroll
        for (int u = 0; u < KL_UNR; ++u) atomicAdd(hist + bb[u] * 32, 1u);
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
    const bool fits = takeall || (counted && cb <= KL_BND);  // else: a boundary bin too full for the list fails below
    const int beq = (takeall || !fits) ? -1 : bstar;
    const int tb = fits ? ((takeall ? KL_NB : bstar + 1) + base) << 21 : 0;     // float bits of the upper edge of b*
    // pass C: the histogram guarantees c0 + cb < K + KL_BND <= KL_LIST survivors, so the list cannot overflow; every
    // rejected d2 is >= tau (the upper edge of b*), which is all the certification needs to know about them
    const unsigned la0 = (unsigned)__cvta_generic_to_shared(list);   // 32-bit shared address: one add per survivor
    unsigned la = la0;
    for (int s = 0; s < np; s += KL_UNR) {
        float4 cc[KL_UNR];
        float dd[KL_UNR];
#pragma unroll
        for (int u = 0; u < KL_UNR; ++u) cc[u] = cd[s + u];
#pragma unroll
        for (int u = 0; u < KL_UNR; u += 2) {
            float2 d = kl_d2pair(cc[u], cc[u + 1], nqx, nqy, nqz);
            dd[u] = d.x; dd[u + 1] = d.y;
        }
#pragma unroll
        for (int u = 0; u < KL_UNR; ++u)
            if (__float_as_int(dd[u]) < tb) {
                kl_sts16(la, (unsigned short)(s + u));
                la += 64;
            }
    }
    const int live = (int)(la - la0) >> 6;                    // = c0 + cb (takeall: everything inside the range)
    const int cmax = __reduce_max_sync(FULL, (unsigned)live);
    __syncwarp();     // the boundary lists alias the histograms of other lanes
    // pass F
    float *bd = sm.u.bnd.d + lane;
    unsigned short *bs = sm.u.bnd.s + lane;
    unsigned bda = (unsigned)__cvta_generic_to_shared(bd), bsa = (unsigned)__cvta_generic_to_shared(bs), nba = la0;
    for (int e = 0; e < cmax; ++e) {
        const int sl = e < live ? (int)list[e * 32] : 0;
        const float d = kl_d2slot(cd, sl, qx, qy, qz);
        const int b = e < live ? kl_bin(d, base) : KL_NB + 1;
        if (b < bstar) { kl_sts16(nba, (unsigned short)sl); nba += 64; }
        else if (b == beq) { kl_sts32(bda, d); kl_sts16(bsa, (unsigned short)sl); bda += 128; bsa += 64; }
    }
    const float ubound = U + kf_err(U, eA, eC) * 1.0001f;
    if (dead) { A.why = 4; A.capf = ubound; }
    else if (takeall) {
        A.ok = tcap + kf_err(tcap, eA, eC) <= rho1f;
        A.capf = tcap;
        if (!A.ok) A.why = 2;
    } else if (!counted) { A.why = 4; A.capf = ubound; }
    else if (cb > KL_BND) {
        A.why = 3;
        float edge = __int_as_float((bstar + base + 1) << 21);      // every d2 of bins <= b* is below it
        A.capf = fminf(ubound, edge + kf_err(edge, eA, eC) * 1.0001f);
    } else {
        const int need = K - c0;
        float a = 0.f;
        for (int r = 0; r < need; ++r) {
            float best = KF_HUGE;
            int be = 0;
            for (int e = 0; e < cb; ++e) { float v = bd[e * 32]; if (v < best) { best = v; be = e; } }
            a = best;
            kl_sts16(nba, bs[be * 32]);
            nba += 64;
            bd[be * 32] = KF_HUGE;
        }
        float rest = __int_as_float(tb);                            // tau: no rejected candidate is closer
        for (int e = 0; e < cb; ++e) rest = fminf(rest, bd[e * 32]);
        const float ea = kf_err(a, eA, eC);
        const bool c1 = rest - a > 2.0f * kf_err(rest, eA, eC), c2 = a + ea <= rho1f;
        A.ok = c1 && c2;
        A.capf = a + ea * 1.0001f;
        if (!c1) A.why = 1; else if (!c2) A.why = 2;
    }
    A.cn = (int)(nba - la0) >> 6;
    return A;
}


// stats slots (out.stats, O3DPP_STATS=1): 16 warps, 17 candidates, 18 groups scanned, 19 fail-gap,
// 20 fail-rho, 21 fail-boundary-overflow, 22 fail-count, 23 fail-not-staged, 24 restages, 25 parts
template <int MODE>
__global__ void __launch_bounds__(KL_WARPS * 32, KL_MINB)
knn_lane_kernel(const GridDesc g, int K, double radius, float rho_scale, int max_shrink, int gsplit, float shrink, KnnOut out,
                float *__restrict__ capbuf)
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

    const long long t_start = out.stats ? clock64() : 0;
    KlStage sg(g, sm);
    sg.lane = lane; sg.scanned = 0; sg.home = leaf_id; sg.gsplit = gsplit;
    // the home box; candidate coordinates are relative to its centre for every part of the group
    double hlo[3], hhi[3];
    {
        const double *b0 = g.box + g.off[0];
        int64_t p0 = g.pitch[0];
        hlo[0] = b0[leaf_id]; hlo[1] = b0[p0 + leaf_id]; hlo[2] = b0[2 * p0 + leaf_id];
        hhi[0] = b0[3 * p0 + leaf_id]; hhi[1] = b0[4 * p0 + leaf_id]; hhi[2] = b0[5 * p0 + leaf_id];
    }
    sg.cc[0] = 0.5 * (hlo[0] + hhi[0]); sg.cc[1] = 0.5 * (hlo[1] + hhi[1]); sg.cc[2] = 0.5 * (hlo[2] + hhi[2]);
    const double hext = fmax(fmax(hhi[0] - hlo[0], hhi[1] - hlo[1]), hhi[2] - hlo[2]);
    // own query in local FP32; the home group occupies slots 0..31 (bit-identical to the query copies)
    const float qx = valid ? (float)(px - sg.cc[0]) : KF_BIG, qy = valid ? (float)(py - sg.cc[1]) : KF_BIG,
                qz = valid ? (float)(pz - sg.cc[2]) : KF_BIG;
    kl_put(sm.cand, lane, qx, qy, qz, valid ? (int)mypos : -1);
    __syncwarp();
    // U = largest d2 to a member of the home group: bounds the k-th d2 when the group holds >= K points
    float U = 0.f;
    {
        const float2 nqx = make_float2(-qx, -qx), nqy = make_float2(-qy, -qy), nqz = make_float2(-qz, -qz);
#pragma unroll 8
        for (int s = 0; s < 32; s += 2) {
            float2 d = kl_d2pair(sm.cand[s], sm.cand[s + 1], nqx, nqy, nqz);
            if (s < nq) U = fmaxf(U, d.x);
            if (s + 1 < nq) U = fmaxf(U, d.y);
        }
    }
    unsigned restage = 0, parts = 0, candsum = 0;
    unsigned n1 = 0, n2 = 0, n3 = 0, n4 = 0, n5 = 0;

    // parts of the group: [start, start + len), len a power of two; a part whose grown box overflows the
    // staging area is halved (mixed-density groups), so every block of lanes is staged at most once
    int start = 0, len = 32;
    while (start < nq) {
        const bool inpart = lane >= start && lane < start + len;
        if (len == 32) {
#pragma unroll
            for (int a = 0; a < 3; ++a) { sg.blo[a] = hlo[a]; sg.bhi[a] = hhi[a]; }
        } else {
            const double INF = pos_inf();
            const bool v = inpart && valid;
            sg.blo[0] = warp_min_d(v ? px : INF); sg.blo[1] = warp_min_d(v ? py : INF); sg.blo[2] = warp_min_d(v ? pz : INF);
            sg.bhi[0] = warp_max_d(v ? px : -INF); sg.bhi[1] = warp_max_d(v ? py : -INF); sg.bhi[2] = warp_max_d(v ? pz : -INF);
        }
        const double ex = sg.bhi[0] - sg.blo[0], ey = sg.bhi[1] - sg.blo[1], ez = sg.bhi[2] - sg.blo[2];
        double rho = sqrt((ex * ex + ey * ey) + ez * ez) * (double)rho_scale;
        if (len < 32) rho *= 1.25;          // smaller parts: the k-th distance is larger relative to the box
        if (MODE == 2 && radius * 1.001 < rho) rho = radius * 1.001;
        bool staged = true;
        for (int attempt = 0;; ++attempt) {
            sg.rho = rho; sg.rho2 = rho * rho * (1.0 + 1e-12);
            if (sg.stage()) break;
            restage++;
            // a smaller radius: the sparse queries of the part then go to tier 2, the dense ones stay, and the retry
            // costs one cheap table walk.  A part that still does not fit after max_shrink retries has a box that is
            // itself too large (sparse points spanning dense structures): it is halved if max_shrink < 12, else all
            // its queries go on to tiers 2 / 3 (measured: cheaper than the extra staging passes of the halves)
            if (attempt >= (len > 4 ? max_shrink : 12)) { staged = false; break; }
            rho *= (double)shrink;
        }
        if (!staged && len > 4 && max_shrink < 12) { len >>= 1; __syncwarp(); continue; }
        parts++;
        // a part that cannot be staged (its own box spans dense structures): stage the KL_WIN cell groups before and
        // after the home group along the curve, unfiltered.  Nothing can be certified against them, but the k-th d2
        // among them is a tight upper bound (tools/knn_stage_study.py: 1.17x the true k-th distance in the median),
        // which lets tier 2 answer these queries instead of the exact search of tier 3.
        bool window = false;
        double wmax = 0.0;
        if (!staged && nq >= K) {
            window = true;
            sg.n = 32;
#pragma unroll
            for (int dlt = -KL_WIN; dlt <= KL_WIN; ++dlt) {
                const int64_t leaf = leaf_id + dlt, p = leaf * 32 + lane;
                const bool v = dlt != 0 && leaf >= 0 && p < g.n;
                double lx = 0, ly = 0, lz = 0;
                if (v) { lx = g.sx[p] - sg.cc[0]; ly = g.sy[p] - sg.cc[1]; lz = g.sz[p] - sg.cc[2]; }
                const unsigned m = __ballot_sync(FULL, v);
                if (v) {
                    kl_put(sm.cand, sg.n + __popc(m & ((1u << lane) - 1u)), (float)lx, (float)ly, (float)lz, (int)p);
                    wmax = fmax(wmax, fmax(fmax(fabs(lx), fabs(ly)), fabs(lz)));
                }
                sg.n += __popc(m);
            }
            wmax = warp_max_d(wmax);
        }
        const int ncand = (staged || window) ? sg.n : 32;
        candsum += ncand;
        const int np = (ncand + KL_UNR - 1) & ~(KL_UNR - 1);
        if (lane < np - ncand) kl_put(sm.cand, ncand + lane, KF_BIG, KF_BIG, KF_BIG, -1);
        __syncwarp();
        // error model of the staged FP32 coordinates: |coordinate| <= half home extent + rho (window: the largest staged)
        const float Lf = (float)((window ? fmax(0.5 * hext, wmax) : 0.5 * hext + rho) * 1.001);
        const float eA = 16.0f * KF_U * Lf, eC = 2.3e-13f * Lf * Lf;
        const float rho1f = staged ? __double2float_rd(rho * rho * (1.0 - 1e-6)) : -1.0f;
        float tcap = KF_TCAP;
        if (MODE == 2) {
            float r2f = __double2float_ru(r2d);
            tcap = r2f * (1.0f + 1e-6f) + 2.0f * kf_err(r2f, eA, eC);
            if (!(tcap < KF_TCAP)) staged = false;
        }
        KlAnswer A;
        A.ok = false; A.cn = 0; A.why = 5; A.capf = __int_as_float(0x7f800000);
        if (nq >= K && (staged || window)) {
            A = kl_select<MODE>(sm, lane, np, K, qx, qy, qz, U, tcap, rho1f, eA, eC);
            if (window) { A.ok = false; A.why = 5; }      // rho1f < 0 already refuses; the bound A.capf is what counts
        } else if (nq >= K) A.capf = U + kf_err(U, eA, eC) * 1.0001f;
        if (out.stats && inpart && valid) { n1 += A.why == 1; n2 += A.why == 2; n3 += A.why == 3; n4 += A.why == 4; n5 += A.why == 5; }
        // OUTPUT: sorted positions of the selected neighbours, [j][query position]; KNN_NONE beyond the count
        const bool put = inpart && valid && A.ok && out.nbr != nullptr;
        if (__any_sync(FULL, put)) {
            uint32_t *dst = out.nbr + mypos;
            for (int j = 0; j < K; ++j)
                if (put) dst[(int64_t)j * out.nbr_pitch] = j < A.cn ? (uint32_t)kl_pos(sm.cand, sm.v.nb[j * 32 + lane]) : KNN_NONE;
        }
        if (inpart && valid) {
            capbuf[mypos] = A.ok ? -fmaxf(A.capf, 1e-30f) : A.capf;   // negative = answered, magnitude bounds the k-th d2
            if (out.keep && A.ok) out.tier[mypos] = 1;
        }
        __syncwarp();
        start += len;
        len = start & -start;
    }
    if (out.stats) {
        unsigned s1 = __reduce_add_sync(FULL, n1), s2 = __reduce_add_sync(FULL, n2), s3 = __reduce_add_sync(FULL, n3),
                 s4 = __reduce_add_sync(FULL, n4), s5 = __reduce_add_sync(FULL, n5);
        if (lane == 0) {
            atomicAdd(out.stats + 16, 1ull);
            atomicAdd(out.stats + 17, (unsigned long long)candsum);
            atomicAdd(out.stats + 18, (unsigned long long)sg.scanned);
            atomicAdd(out.stats + 19, (unsigned long long)s1);
            atomicAdd(out.stats + 20, (unsigned long long)s2);
            atomicAdd(out.stats + 21, (unsigned long long)s3);
            atomicAdd(out.stats + 22, (unsigned long long)s4);
            atomicAdd(out.stats + 23, (unsigned long long)s5);
            atomicAdd(out.stats + 24, (unsigned long long)restage);
            atomicAdd(out.stats + 25, (unsigned long long)parts);
            atomicAdd(out.stats + (parts > 1 ? 27 : 26), (unsigned long long)(clock64() - t_start));
            if (parts > 1) atomicAdd(out.stats + 28, 1ull);
        }
    }
}


// ---- tier 2: warp per failed query, exact FP64 range query within cap --------------------------------
// A warp takes 32 consecutive entries of the compacted failure list.  For each it walks the range tables
// with the fixed bound cap (lane = child box / lane = point, coalesced), appends every point with
// d2 <= cap to a shared list, ranks the list by (d2, original index) and sums the cumulants of the K
// smallest; the totals are parked in lane i, and at the end every lane finishes one query.
#ifndef K2_WARPS
#define K2_WARPS 4
#endif
#ifndef K2_SHRINK
#define K2_SHRINK 0     // 1: a full list keeps its K smallest and lowers the bound; 0: the query goes to tier 3
#endif
#ifndef K2_LIST
#define K2_LIST 128
#endif

struct K2WarpMem {
    double d[K2_LIST], x[K2_LIST], y[K2_LIST], z[K2_LIST];
    unsigned oidx[K2_LIST];
    unsigned pos[K2_LIST];      // sorted position (neighbour-set record, keep mode)
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

template <int MODE>
struct K2Range {
    const GridDesc &g;
    K2WarpMem &sm;
    int lane;
    double qx, qy, qz, cap, r2;
    float fql[3], fqh[3], capf;     // the query rounded down / up and the bound rounded up, for the FP32 box test
    int cnt, K;
    __device__ __forceinline__ K2Range(const GridDesc &g_, K2WarpMem &sm_) : g(g_), sm(sm_) {}

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
        if (hit && at < K2_LIST) { sm.d[at] = d; sm.x[at] = x; sm.y[at] = y; sm.z[at] = z; sm.oidx[at] = g.sidx[p]; sm.pos[at] = (unsigned)p; }
        cnt += __popc(m);
        if (K2_SHRINK && cnt > K2_LIST - 32) shrink();
    }

    // the list is nearly full: keep its K smallest by (d2, original index) and lower the bound to the K-th
    __device__ __forceinline__ void shrink()
    {
        __syncwarp();
        constexpr int PER = K2_LIST / 32;
        double de[PER], xe[PER], ye[PER], ze[PER];
        unsigned oe[PER], pe[PER];
        int rank[PER];
#pragma unroll
        for (int a = 0; a < PER; ++a) {
            int e = a * 32 + lane;
            bool ve = e < cnt;
            de[a] = ve ? sm.d[e] : pos_inf(); oe[a] = ve ? sm.oidx[e] : 0xffffffffu; pe[a] = ve ? sm.pos[e] : 0u;
            xe[a] = ve ? sm.x[e] : 0.0; ye[a] = ve ? sm.y[e] : 0.0; ze[a] = ve ? sm.z[e] : 0.0;
            rank[a] = 0;
        }
        for (int f = 0; f < cnt; ++f) {
            double df = sm.d[f];
            unsigned of = sm.oidx[f];
#pragma unroll
            for (int a = 0; a < PER; ++a) rank[a] += (df < de[a] || (df == de[a] && of < oe[a])) ? 1 : 0;
        }
        __syncwarp();
        double kth = 0.0;
#pragma unroll
        for (int a = 0; a < PER; ++a) {
            bool keep = a * 32 + lane < cnt && rank[a] < K;      // ranks are distinct: exactly K survive
            if (keep) { int r = rank[a]; sm.d[r] = de[a]; sm.x[r] = xe[a]; sm.y[r] = ye[a]; sm.z[r] = ze[a]; sm.oidx[r] = oe[a]; sm.pos[r] = pe[a]; }
            if (keep && rank[a] == K - 1) kth = de[a];
        }
#pragma unroll
        for (int dlt = 16; dlt > 0; dlt >>= 1) kth = fmax(kth, shfl_xor_d(kth, dlt));
        cap = kth;
        capf = __double2float_ru(kth);
        cnt = K;
        __syncwarp();
    }

    template <int LVL>
    __device__ __forceinline__ void visit(int64_t parent)
    {
        int64_t c = parent * 32 + lane;
        bool ex = c < g.cnt[LVL];
        float lb = ex ? kl_box_lb(g, LVL, c, fql[0], fqh[0], fql[1], fqh[1], fql[2], fqh[2]) : KF_HUGE;
        unsigned m = __ballot_sync(FULL, ex && lb <= capf);
        if constexpr (!K2_SHRINK) {
            // fixed bound: every child that intersects the ball is opened, the order does not matter
            while (m && cnt <= K2_LIST) {
                const int b = __ffs(m) - 1;
                m &= m - 1;
                if constexpr (LVL == 0) leaf(parent * 32 + b);
                else visit<LVL - 1>(parent * 32 + b);
            }
        } else {
            // nearest child first: the bound drops early (shrink) and the far children are never opened
            const unsigned key = __float_as_uint(lb);
            while (m) {
                m &= __ballot_sync(FULL, lb <= capf);     // the bound may have dropped
                if (!m) break;
                const bool mine = (m >> lane) & 1u;
                const unsigned kmin = __reduce_min_sync(FULL, mine ? key : 0xffffffffu);
                int b = __ffs(__ballot_sync(FULL, mine && key == kmin)) - 1;
                m &= ~(1u << b);
                if constexpr (LVL == 0) leaf(parent * 32 + b);
                else visit<LVL - 1>(parent * 32 + b);
            }
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
                 float *__restrict__ capbuf)
{
    __shared__ K2WarpMem k2_sm[K2_WARPS];
    const int lane = lane_id(), w = threadIdx.x >> 5;
    const int64_t first = ((int64_t)blockIdx.x * K2_WARPS + w) * 32;
    if (first >= m) return;
    const int nq = (int)(m - first < 32 ? m - first : 32);
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
    K2Range<MODE> rq(g, sm);
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
        if (cnt > K2_LIST) continue;             // (K2_SHRINK == 0) too many candidates inside the bound: tier 3
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
