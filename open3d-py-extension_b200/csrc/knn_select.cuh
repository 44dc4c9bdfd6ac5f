// knn_select.cuh — tier 1 of estimate_normals (reference call: PointCloud.py:68-73 -> KDTreeFlann::SearchKNN / SearchHybrid
// inside Open3D's EstimateNormals): certified FP32 selection of the K nearest neighbours of every point, as two kernels.
//
// knn_groups_kernel (one warp per cell group = 32 Hilbert-consecutive points; full occupancy):
//   walks the range tables level by level and lists the cell groups whose box comes within rho = rho_scale x diagonal of
//   the home box.  A list longer than gsplit retries with the radius x shrink (the group's dense queries still certify,
//   its sparse ones go to tier 2).  Output: the list, its length, the retries used.
// knn_select_kernel (KS_W warps cooperate on ONE cell group = 32 queries; bound by shared memory, ~19 KB per group):
//   STAGE   the listed groups are loaded KS_SB at a time, alternately by the warps; points inside the home box grown by
//           rho are appended to shared memory as FP32 coordinates relative to the home-box centre (recentred in FP64),
//           two candidates per 32-byte record {x0 x1 y0 y1 | z0 z1 p0 p1}.  Warp 0 fills the staging area upwards,
//           warp 1 downwards, so the slot order is fixed whatever the timing; if the two regions meet, the area is too
//           small: the radius shrinks, the list is filtered in place (a smaller rho lists a subset), and staging repeats.
//   SELECT  lane = query in BOTH warps; each warp walks its own region of candidates for all 32 queries in lockstep (a
//           candidate is a broadcast LDS.128, no divergence):
//             pass A  d2 of every candidate -> shared histogram [bin][lane] of 32 quarter-octave bins below U (the largest
//                     d2 to a member of the home group, which bounds the k-th d2) by fire-and-forget ATOMS;
//             scan    b* = the bin in which the cumulative count crosses K (both warps, redundantly);
//             pass C  d2 below the lower edge of b* -> neighbour list, inside b* -> boundary list with its d2; both lists
//                     are lane-private columns filled from the bottom by warp 0 and from the top by warp 1;
//             pick    (warp 0) the K - c0 smallest boundary entries complete the set; certification of the k-th against
//                     the (k+1)-th d2 (gap > 2E, E bounds |d2_fp32 - d2_fp64|) and against rho (complete);
//   OUTPUT  the K sorted positions of every certified lane -> global [j][query position] (coalesced), both warps;
//           lanes that cannot certify publish cap = a + E, an upper bound of the true k-th d2, for tier 2.
// The finish (FP64 cumulants, eigenvector) is knn_finish_kernel.  Every certified set equals the exact FP64 (d2, index)
// set, so the result does not depend on KS_W; the neighbour ORDER (summation order of the cumulants) is fixed per KS_W.
#pragma once
#include "knn_finish.cuh"

#ifndef KS_W
#define KS_W 1          // warps cooperating on one cell group (1 or 2; measured at 10 M: 1 -> 6.58 ms, 2 -> 7.69 ms: the barriers and the work both warps repeat cost more than the shorter critical path saves)
#endif
#ifndef KS_CAP
#define KS_CAP 768      // staging slots per group, 2 KS_UNR of them padding
#endif
#ifndef KS_NB
#define KS_NB 32        // histogram bins
#endif
#ifndef KS_SHIFT
#define KS_SHIFT 21     // bin = float bits of d2 >> KS_SHIFT: quarter octaves, 8 octaves in 32 bins (20: boundary fails 0.8 % -> 13.6 %)
#endif
#ifndef KS_UNR
#define KS_UNR 8        // candidates per iteration of the select passes
#endif
#ifndef KS_SB
#define KS_SB 4         // cell groups loaded per staging batch (3 KS_SB loads in flight per lane)
#endif
#ifndef KS_BND
#define KS_BND 22       // boundary-bin entries per query
#endif
#ifndef KS_GMAX
#define KS_GMAX 512     // frontier entries of the table walk; upper bound of gsplit + 64
#endif
#ifndef KS_BFB
#define KS_BFB 2        // table walk: frontier nodes whose child boxes are loaded together
#endif
#ifndef KS_WIN
#define KS_WIN 2        // unstageable groups: cell groups taken before and after the home group along the curve
#endif
#ifndef KS_LANE_RHO
#define KS_LANE_RHO 1   // certify completeness against rho + the query's own clearance inside the home box instead of rho
#endif
#ifndef KS_MINB
#define KS_MINB (KS_W == 2 ? 10 : 11)    // resident blocks per SM the register budget is set for
#endif
#ifndef KS_STAGE4
#define KS_STAGE4 0     // 1: staging reads the 32-byte point records instead of the three coordinate planes (fewer instructions, a third more bytes: 6.62 ms against 6.52 ms at 10 M)
#endif
#ifndef KA_WARPS
#define KA_WARPS 8
#endif
#ifndef KS_LIST
#define KS_LIST 55      // survivors of the threshold pass per query (16-bit staged slots): >= K - 1 + KS_BND + 1; 55 keeps the block at 20 064 B = 11 groups per SM
#endif
#define KS_NOLIST 0xffffu

static_assert(KS_W == 1 || KS_W == 2, "one or two warps per cell group");
static_assert(32 - 1 + KS_BND < KS_LIST, "survivor list: K - 1 below the boundary bin + KS_BND inside it (K <= 32 on this path)");
static_assert(KS_CAP % KS_UNR == 0 && KS_UNR % 2 == 0, "regions are padded to whole select iterations");

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
// bin of a non-negative float relative to `base`: monotone in d2, three integer instructions; KS_NB = beyond the range
__device__ __forceinline__ int kl_bin(float d, int base)
{
    return __vimin_s32_relu((__float_as_int(d) >> KS_SHIFT) - base, KS_NB);      // max(min(b, KS_NB), 0) in one instruction
}

// lower bound of the squared gap between the home box (flo, fhi: rounded outwards) and box c of table lvl, in FP32 with
// every operation rounded down and the stored boxes rounded outwards: a group that matters is never culled
__device__ __forceinline__ float ks_gap(const GridDesc &g, int lvl, int64_t c, const float *flo, const float *fhi)
{
    const float4 *e = g.boxq + 2 * (g.qoff[lvl] + (int)c);
    const float4 a = e[0];
    const float2 b = *(const float2 *)(e + 1);
    float gx = fmaxf(fmaxf(__fsub_rd(a.x, fhi[0]), __fsub_rd(flo[0], a.w)), 0.f);
    float gy = fmaxf(fmaxf(__fsub_rd(a.y, fhi[1]), __fsub_rd(flo[1], b.x)), 0.f);
    float gz = fmaxf(fmaxf(__fsub_rd(a.z, fhi[2]), __fsub_rd(flo[2], b.y)), 0.f);
    return __fadd_rd(__fadd_rd(__fmul_rd(gx, gx), __fmul_rd(gy, gy)), __fmul_rd(gz, gz));
}

struct KsHome {                 // what both kernels derive from the home box, identically
    double lo[3], hi[3], cc[3], hext, rho0;
    float flo[3], fhi[3];
};
template <int MODE>
__device__ __forceinline__ void ks_home(const GridDesc &g, int64_t leaf_id, double radius, float rho_scale, KsHome &h)
{
    const double *b0 = g.box + g.off[0];
    const int64_t p0 = g.pitch[0];
#pragma unroll
    for (int a = 0; a < 3; ++a) { h.lo[a] = b0[a * p0 + leaf_id]; h.hi[a] = b0[(3 + a) * p0 + leaf_id]; }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        h.cc[a] = 0.5 * (h.lo[a] + h.hi[a]);
        h.flo[a] = __double2float_rd(h.lo[a]); h.fhi[a] = __double2float_ru(h.hi[a]);
    }
    const double ex = h.hi[0] - h.lo[0], ey = h.hi[1] - h.lo[1], ez = h.hi[2] - h.lo[2];
    h.hext = fmax(fmax(ex, ey), ez);
    double rho = sqrt((ex * ex + ey * ey) + ez * ez) * (double)rho_scale;
    if (MODE == 2 && radius * 1.001 < rho) rho = radius * 1.001;
    h.rho0 = rho;
}

struct KsGroups {               // knn_groups_kernel -> knn_select_kernel
    uint32_t *list;             // [group][pitch] cell groups to scan, ascending
    unsigned short *cnt;        // entries, KS_NOLIST = no radius with a short enough list was found
    unsigned char *tries;       // shrink retries behind the list: rho = rho0 x shrink^tries
    int pitch;
};

// ---- the group lists ------------------------------------------------------------------------------------------------
// Level-by-level walk: the frontier (the nodes of one level whose box comes within rho) ping-pongs between two shared
// arrays; the child boxes of KS_BFB frontier nodes are loaded together through a clamped index, so a level costs one
// memory latency instead of one per node.  Children are appended in node order: the list is ascending.
template <int MODE>
__global__ void __launch_bounds__(KA_WARPS * 32)
knn_groups_kernel(const GridDesc g, double radius, float rho_scale, int max_shrink, int gsplit, float shrink, KsGroups gr)
{
    __shared__ unsigned ka_buf[KA_WARPS][2][KS_GMAX];
    const int lane = lane_id(), w = threadIdx.x >> 5;
    const int64_t leaf_id = (int64_t)blockIdx.x * KA_WARPS + w;
    if (leaf_id >= g.cnt[0]) return;
    if (g.qfirst < g.n) {                        // O3DPP_OPT_QUERY_FIRST: a group none of whose points is a query needs no list
        const int64_t p = leaf_id * 32 + lane;
        if (!__any_sync(FULL, p < g.n && (int64_t)g.sidx[p] < g.qfirst)) return;
    }
    KsHome h;
    ks_home<MODE>(g, leaf_id, radius, rho_scale, h);
    const unsigned lt = (1u << lane) - 1u;
    const int top = g.levels - 1;
    double rho = h.rho0;
    int ngroups = -1, attempt = 0;
    for (;; ++attempt) {
        const float rho2f = __double2float_ru(rho * rho * (1.0 + 1e-12));
        __syncwarp();
        if (lane == 0) ka_buf[w][(top + 1) & 1][0] = 0u;
        int ncur = 1;
        bool over = false;
        for (int lvl = top; lvl >= 0 && !over; --lvl) {
            __syncwarp();
            const unsigned *cur = ka_buf[w][(lvl + 1) & 1];
            unsigned *nxt = ka_buf[w][lvl & 1];
            const int64_t cnt = g.cnt[lvl];
            const int lim = lvl ? KS_GMAX - 32 * KS_BFB : gsplit;
            int nout = 0, i = 0;
#if KS_BFB > 1
            for (; i + KS_BFB <= ncur && nout <= lim; i += KS_BFB) {
                int64_t c[KS_BFB];
                float gp[KS_BFB];
#pragma unroll
                for (int k = 0; k < KS_BFB; ++k) {
                    c[k] = (int64_t)cur[i + k] * 32 + lane;
                    gp[k] = ks_gap(g, lvl, c[k] < cnt ? c[k] : cnt - 1, h.flo, h.fhi);
                }
#pragma unroll
                for (int k = 0; k < KS_BFB; ++k) {
                    const bool take = c[k] < cnt && gp[k] <= rho2f && (lvl > 0 || c[k] != leaf_id);
                    const unsigned m = __ballot_sync(FULL, take);
                    const int at = nout + __popc(m & lt);
                    if (take && at < KS_GMAX) nxt[at] = (unsigned)c[k];
                    nout += __popc(m);
                }
            }
#endif
#pragma unroll 1
            for (; i < ncur && nout <= lim; ++i) {
                const int64_t c = (int64_t)cur[i] * 32 + lane;
                const float gp = ks_gap(g, lvl, c < cnt ? c : cnt - 1, h.flo, h.fhi);
                const bool take = c < cnt && gp <= rho2f && (lvl > 0 || c != leaf_id);
                const unsigned m = __ballot_sync(FULL, take);
                const int at = nout + __popc(m & lt);
                if (take && at < KS_GMAX) nxt[at] = (unsigned)c;
                nout += __popc(m);
            }
            if (nout > lim) over = true;       // more groups than gsplit (or a frontier that cannot be held): shrink
            ncur = nout;
        }
        if (!over) { ngroups = ncur; break; }
        if (attempt >= max_shrink) break;
        rho *= (double)shrink;
    }
    __syncwarp();
    if (ngroups >= 0) {
        uint32_t *dst = gr.list + leaf_id * gr.pitch;
        for (int i = lane; i < ngroups; i += 32) dst[i] = ka_buf[w][0][i];
    }
    if (lane == 0) {
        gr.cnt[leaf_id] = ngroups >= 0 ? (unsigned short)ngroups : (unsigned short)KS_NOLIST;
        gr.tries[leaf_id] = (unsigned char)attempt;
    }
}

// ---- the selection ---------------------------------------------------------------------------------------------------
struct KsMem {
    float4 cand[KS_CAP];
    union {
        unsigned hist[(KS_NB + 1) * 32];                                          // [bin][lane]
        struct { float d[KS_BND * 32]; unsigned short s[KS_BND * 32]; } bnd;      // [entry][lane], two-ended
    } u;
    union {
        unsigned short nb[KS_LIST * 32];                                          // [j][lane] staged slots: survivors, then the neighbours; two-ended
        unsigned gl[KS_GMAX];                                                     // cell groups to scan (staging only)
    } v;
#if KS_W > 1                                  // what the two warps tell each other (one warp: registers; 20.1 KB per group = 11 groups per SM)
    unsigned short na[2][32], nbd[2][32];     // pass F, per warp and lane: neighbours / boundary entries collected
    unsigned char ok[32], cn_up[32], cn_dn[32];
#endif
    int n_lo, n_hi, ng;
    double wmax;
};

template <int W>
__device__ __forceinline__ void ks_sync()
{
    if (W == 1) __syncwarp();
    else __syncthreads();
}

// stats slots (out.stats, O3DPP_OPT_STATS): 16 groups, 17 candidates, 18 groups scanned, 19 fail-gap, 20 fail-rho,
// 21 fail-boundary-overflow, 22 fail-count, 23 fail-not-staged, 24 restages
template <int MODE, int W>
__global__ void __launch_bounds__(W * 32, KS_MINB)
knn_select_kernel(const GridDesc g, int K, double radius, float rho_scale, int max_shrink, float shrink, const KsGroups gr,
                  KnnOut out, float *__restrict__ capbuf)
{
    extern __shared__ __align__(16) unsigned char ks_smem[];
    KsMem &sm = *(KsMem *)ks_smem;
    const int lane = lane_id(), w = W > 1 ? (int)(threadIdx.x >> 5) : 0;
    const unsigned lt = (1u << lane) - 1u;
    const int64_t leaf_id = blockIdx.x;
    const int64_t base = leaf_id * 32;
    const int nq = (int)(g.n - base < 32 ? g.n - base : 32);
    const int64_t mypos = base + lane;
    const bool valid = lane < nq;
    double px = 0, py = 0, pz = 0;
    bool want = valid;                           // O3DPP_OPT_QUERY_FIRST: only the points with original index < qfirst are queries
    if (g.qfirst < g.n) {
        want = valid && (int64_t)g.sidx[mypos] < g.qfirst;
        if (!__any_sync(FULL, want)) {           // nothing to answer here (both warps of a pair agree): the whole block leaves
            if (w == 0 && valid) capbuf[mypos] = KF_SKIP;
            return;
        }
    }
    if (valid) { px = g.sx[mypos]; py = g.sy[mypos]; pz = g.sz[mypos]; }
    KsHome h;
    ks_home<MODE>(g, leaf_id, radius, rho_scale, h);
    // own query in local FP32; the home group occupies slots 0..31 (bit-identical to the query copies)
    const float qx = valid ? (float)(px - h.cc[0]) : KF_BIG, qy = valid ? (float)(py - h.cc[1]) : KF_BIG,
                qz = valid ? (float)(pz - h.cc[2]) : KF_BIG;
    if (w == 0) kl_put(sm.cand, lane, qx, qy, qz, valid ? (int)mypos : -1);
    int ng = gr.cnt[leaf_id], tries = gr.tries[leaf_id];
    bool staged = ng != (int)KS_NOLIST;
    if (staged) {
        const uint32_t *src = gr.list + leaf_id * gr.pitch;
        for (int i = threadIdx.x; i < ng; i += W * 32) sm.v.gl[i] = src[i];
    }
    ks_sync<W>();
    const float2 nqx = make_float2(-qx, -qx), nqy = make_float2(-qy, -qy), nqz = make_float2(-qz, -qz);
    // U = largest d2 to a member of the home group: bounds the k-th d2 when the group holds >= K points
    float U = 0.f;
#pragma unroll 8
    for (int s = 0; s < 32; s += 2) {
        float2 d = kl_d2pair(sm.cand[s], sm.cand[s + 1], nqx, nqy, nqz);
        if (s < nq) U = fmaxf(U, d.x);
        if (s + 1 < nq) U = fmaxf(U, d.y);
    }
    double rho = h.rho0;
    for (int t = 0; t < tries; ++t) rho *= (double)shrink;
    unsigned restage = 0, scanned = 0;

    // STAGE: the points of the listed groups that fall inside the home box grown by rho
    int n_lo = 32, n_hi = 0;
    const int64_t last = g.n - 1;
    while (staged) {
        const double xl = h.lo[0] - rho, xh = h.hi[0] + rho, yl = h.lo[1] - rho, yh = h.hi[1] + rho, zl = h.lo[2] - rho, zh = h.hi[2] + rho;
        int n = w == 0 ? 32 : 0;                 // entries of this warp's region (warp 0: the home group first)
        for (int i = w * KS_SB; i < ng && n <= KS_CAP; i += W * KS_SB) {
            int64_t p[KS_SB];
            double x[KS_SB], y[KS_SB], z[KS_SB];
#pragma unroll
            for (int k = 0; k < KS_SB; ++k) {
                p[k] = i + k < ng ? (int64_t)sm.v.gl[i + k] * 32 + lane : g.n;
                const int64_t q = p[k] < last ? p[k] : last;
#if KS_STAGE4
                const double4 c4 = g.sp4[q];
                x[k] = c4.x; y[k] = c4.y; z[k] = c4.z;
#else
                x[k] = g.sx[q]; y[k] = g.sy[q]; z[k] = g.sz[q];
#endif
            }
#pragma unroll
            for (int k = 0; k < KS_SB; ++k) {
                const bool in = p[k] < g.n && x[k] >= xl && x[k] <= xh && y[k] >= yl && y[k] <= yh && z[k] >= zl && z[k] <= zh;
                const unsigned m = __ballot_sync(FULL, in);
                const int r = n + __popc(m & lt);
                const int at = w == 0 ? r : KS_CAP - 1 - r;
                if (in && r < (w == 0 ? KS_CAP : KS_CAP - 32)) kl_put(sm.cand, at, (float)(x[k] - h.cc[0]), (float)(y[k] - h.cc[1]), (float)(z[k] - h.cc[2]), (int)p[k]);
                n += __popc(m);
            }
            scanned += KS_SB;
        }
        if (W > 1) {
            if (lane == 0) { if (w == 0) sm.n_lo = n; else sm.n_hi = n; }
            __syncthreads();
            n_lo = sm.n_lo; n_hi = sm.n_hi;
        } else n_lo = n;
        if (n_lo + n_hi <= KS_CAP - 2 * KS_UNR) break;
        // the regions met (their content is void): a smaller radius, its list is a subset of this one
        restage++;
        if (tries >= max_shrink) { staged = false; break; }
        tries++;
        rho *= (double)shrink;
        const float rho2f = __double2float_ru(rho * rho * (1.0 + 1e-12));
        if (w == 0) {
            int nout = 0;
            for (int i0 = 0; i0 < ng; i0 += 32) {
                const int i = i0 + lane;
                const unsigned c = i < ng ? sm.v.gl[i] : 0u;
                const bool keep = i < ng && ks_gap(g, 0, c, h.flo, h.fhi) <= rho2f;
                const unsigned m = __ballot_sync(FULL, keep);
                if (keep) sm.v.gl[nout + __popc(m & lt)] = c;
                nout += __popc(m);
                __syncwarp();
            }
            if (lane == 0) sm.ng = nout;
        }
        ks_sync<W>();
        ng = sm.ng;
        ks_sync<W>();
    }
    // a group that cannot be staged (its own box spans dense structures): the KS_WIN cell groups before and after it along
    // the curve, unfiltered.  Nothing can be certified against them, but their k-th d2 is a tight upper bound (1.17x the
    // true k-th distance in the median), which lets tier 2 answer these queries instead of the exact search of tier 3.
    const bool window = !staged && nq >= K;
    double wmax = 0.0;
    if (window) {
        if (w == 0) {
            int n = 32;
#pragma unroll
            for (int dlt = -KS_WIN; dlt <= KS_WIN; ++dlt) {
                const int64_t leaf = leaf_id + dlt, p = leaf * 32 + lane;
                const bool v = dlt != 0 && leaf >= 0 && p < g.n;
                double lx = 0, ly = 0, lz = 0;
                if (v) { lx = g.sx[p] - h.cc[0]; ly = g.sy[p] - h.cc[1]; lz = g.sz[p] - h.cc[2]; }
                const unsigned m = __ballot_sync(FULL, v);
                if (v) {
                    kl_put(sm.cand, n + __popc(m & lt), (float)lx, (float)ly, (float)lz, (int)p);
                    wmax = fmax(wmax, fmax(fmax(fabs(lx), fabs(ly)), fabs(lz)));
                }
                n += __popc(m);
            }
            wmax = warp_max_d(wmax);
            if (lane == 0) { sm.n_lo = n; sm.wmax = wmax; }
        }
        ks_sync<W>();
        n_lo = sm.n_lo; n_hi = 0; wmax = sm.wmax;
    }
    const bool have = staged || window;
    if (!have) { n_lo = 32; n_hi = 0; }
    // regions of the two warps, padded to whole iterations of the select passes
    const int np0 = (n_lo + KS_UNR - 1) & ~(KS_UNR - 1);
    const int s1 = (KS_CAP - n_hi) & ~(KS_UNR - 1);
    if (w == 0) { if (lane < np0 - n_lo) kl_put(sm.cand, n_lo + lane, KF_BIG, KF_BIG, KF_BIG, -1); }
    else if (lane < KS_CAP - n_hi - s1) kl_put(sm.cand, s1 + lane, KF_BIG, KF_BIG, KF_BIG, -1);
    const int sbeg = w == 0 ? 0 : s1, send = w == 0 ? np0 : KS_CAP;
    // error model of the staged FP32 coordinates: |coordinate| <= half home extent + rho (window: the largest staged)
    const float Lf = (float)((window ? fmax(0.5 * h.hext, wmax) : 0.5 * h.hext + rho) * 1.001);
    const float eA = 16.0f * KF_U * Lf, eC = 2.3e-13f * Lf * Lf;
    // completeness radius of THIS lane: every point within rho of the home box is staged (its group's box is within rho of the
    // home box, so the group is listed, and the point lies inside the grown box), and the ball of radius delta = the query's
    // distance to the nearest face of the home box lies inside the home box: everything within rho + delta of the query is staged
    double delta = 0.0;
    if (valid) delta = fmin(fmin(fmin(px - h.lo[0], h.hi[0] - px), fmin(py - h.lo[1], h.hi[1] - py)), fmin(pz - h.lo[2], h.hi[2] - pz));
#if KS_LANE_RHO
    const double rlane = rho + fmax(delta, 0.0) * (1.0 - 1e-9);
#else
    const double rlane = rho;
#endif
    const float rho1f = staged ? __double2float_rd(rlane * rlane * (1.0 - 1e-6)) : -1.0f;
    float tcap = KF_TCAP;
    bool select = nq >= K && have;
    if (MODE == 2) {
        const float r2f = __double2float_ru(__dmul_rn(radius, radius));
        tcap = r2f * (1.0f + 1e-6f) + 2.0f * kf_err(r2f, eA, eC);
        if (!(tcap < KF_TCAP)) select = false;
    }
    // what this lane learns about its query
    bool ok = false;
    int cn_up = 0, cn_dn = 0;
    float capf = __int_as_float(0x7f800000);
    unsigned why = 5;
    if (select) {
        float T = U;
        bool capped = false;
        if (MODE == 2 && tcap < T) { T = tcap; capped = true; }
        const bool dead = !(T > 1e-36f);
        const int bbase = (__float_as_int(dead ? 1.0f : T) >> KS_SHIFT) - (KS_NB - 1);
        unsigned *hist = sm.u.hist + lane;
        for (int b = w; b < KS_NB; b += W) hist[b * 32] = 0u;
        ks_sync<W>();
        const float4 *cd = sm.cand;
        // pass A (candidates are broadcast reads; unconditional shared atomics, bin KS_NB = beyond the range)
        for (int s = sbeg; s < send; s += KS_UNR) {
            float4 cc[KS_UNR];
            int bb[KS_UNR];
#pragma unroll
            for (int u = 0; u < KS_UNR; ++u) cc[u] = cd[s + u];
#pragma unroll
            for (int u = 0; u < KS_UNR; u += 2) {
                float2 d = kl_d2pair(cc[u], cc[u + 1], nqx, nqy, nqz);
                bb[u] = kl_bin(d.x, bbase); bb[u + 1] = kl_bin(d.y, bbase);
            }
#pragma unroll
            for (int u = 0; u < KS_UNR; ++u) atomicAdd(hist + bb[u] * 32, 1u);
        }
        ks_sync<W>();
        int c0 = 0, bstar = -1, cb = 0, acc = 0;
#pragma unroll
        for (int b = 0; b < KS_NB; ++b) {
            int hv = (int)hist[b * 32];
            if (bstar < 0 && acc + hv >= K) { bstar = b; c0 = acc; cb = hv; }
            acc += hv;
        }
        const bool takeall = bstar < 0 && MODE == 2 && capped;   // fewer than K inside the radius
        const bool counted = bstar >= 0;
        if (takeall) bstar = KS_NB;
        const bool fits = takeall || (counted && cb <= KS_BND);  // else: a boundary bin too full for the list fails below
        const int thi = fits ? ((takeall ? KS_NB : bstar + 1) + bbase) << KS_SHIFT : 0;     // float bits of the upper edge of b*
        ks_sync<W>();     // the boundary lists alias the histograms: every lane of both warps has read its column
        // pass C: d2 below the upper edge of b* -> the slot joins the survivor list: one compare, one predicated store and one
        // predicated add per candidate.  The histogram guarantees c0 + cb < K + KS_BND <= KS_LIST survivors over both warps, so
        // the two ends of the column cannot meet; every rejected d2 is >= the edge, which is all the certification needs
        constexpr int LTOP = KS_LIST - 1;
        const unsigned la0 = (unsigned)__cvta_generic_to_shared(sm.v.nb + (w == 0 ? 0 : LTOP * 32) + lane);
        const unsigned bd0 = (unsigned)__cvta_generic_to_shared(sm.u.bnd.d + (w == 0 ? 0 : (KS_BND - 1) * 32) + lane);
        const unsigned bs0 = (unsigned)__cvta_generic_to_shared(sm.u.bnd.s + (w == 0 ? 0 : (KS_BND - 1) * 32) + lane);
        const int stn = w == 0 ? 64 : -64, std_ = w == 0 ? 128 : -128;
        unsigned la = la0;
        for (int s = sbeg; s < send; s += KS_UNR) {
            float4 cc[KS_UNR];
            float dd[KS_UNR];
#pragma unroll
            for (int u = 0; u < KS_UNR; ++u) cc[u] = cd[s + u];
#pragma unroll
            for (int u = 0; u < KS_UNR; u += 2) {
                float2 d = kl_d2pair(cc[u], cc[u + 1], nqx, nqy, nqz);
                dd[u] = d.x; dd[u + 1] = d.y;
            }
#pragma unroll
            for (int u = 0; u < KS_UNR; ++u)
                if (__float_as_int(dd[u]) < thi) {
                    kl_sts16(la, (unsigned short)(s + u));
                    la += stn;
                }
        }
        // pass F: each warp looks at its own survivors again, compacting towards its own end of the column (the write index never
        // passes the read index): bins below b* are neighbours, those of b* go to the boundary list with their d2
        const int live = (int)(la - la0) / stn;
        const int cmax = __reduce_max_sync(FULL, (unsigned)live);
        const int beq = (takeall || !fits) ? -1 : bstar;
        unsigned nba = la0, bda = bd0, bsa = bs0;
        {
            const unsigned short *list = sm.v.nb + lane;
            for (int e = 0; e < cmax; ++e) {
                const int sl = e < live ? (int)list[(w == 0 ? e : LTOP - e) * 32] : 0;
                const float d = kl_d2slot(cd, sl, qx, qy, qz);
                const int b = e < live ? kl_bin(d, bbase) : KS_NB + 1;
                if (b < bstar) { kl_sts16(nba, (unsigned short)sl); nba += stn; }
                else if (b == beq) { kl_sts32(bda, d); kl_sts16(bsa, (unsigned short)sl); bda += std_; bsa += stn; }
            }
        }
        const int mya = (int)(nba - la0) / stn, myb = (int)(bsa - bs0) / stn;
#if KS_W > 1
        sm.na[w][lane] = (unsigned short)mya;
        sm.nbd[w][lane] = (unsigned short)myb;
#endif
        ks_sync<W>();
        if (w == 0) {
#if KS_W > 1
            const int a0 = sm.na[0][lane], a1 = sm.na[1][lane], b0 = sm.nbd[0][lane], b1 = sm.nbd[1][lane];
#else
            const int a0 = mya, a1 = 0, b0 = myb, b1 = 0;
#endif
            float *bd = sm.u.bnd.d + lane;               // warp 0's entries at [0, b0), warp 1's at [KS_BND - b1, KS_BND)
            const unsigned short *bs = sm.u.bnd.s + lane;
            const float ubound = U + kf_err(U, eA, eC) * 1.0001f;
            cn_up = a0; cn_dn = a1;
            why = 0;
            if (dead) { why = 4; capf = ubound; }
            else if (takeall) {
                ok = tcap + kf_err(tcap, eA, eC) <= rho1f;
                capf = tcap;
                if (!ok) why = 2;
            } else if (!counted) { why = 4; capf = ubound; }
            else if (cb > KS_BND) {
                why = 3;
                const float edge = __int_as_float((bstar + bbase + 1) << KS_SHIFT);      // every d2 of bins <= b* is below it
                capf = fminf(ubound, edge + kf_err(edge, eA, eC) * 1.0001f);
            } else {
                const int need = K - c0;
                unsigned nbw = (unsigned)__cvta_generic_to_shared(sm.v.nb + a0 * 32 + lane);
                float a = 0.f;
                for (int r = 0; r < need; ++r) {
                    float best = KF_HUGE;
                    int be = 0;
                    for (int e = 0; e < b0; ++e) { float v = bd[e * 32]; if (v < best) { best = v; be = e; } }
                    if (W > 1) for (int e = KS_BND - b1; e < KS_BND; ++e) { float v = bd[e * 32]; if (v < best) { best = v; be = e; } }
                    a = best;
                    kl_sts16(nbw, bs[be * 32]);
                    nbw += 64;
                    bd[be * 32] = KF_HUGE;
                }
                cn_up = a0 + need;
                float rest = __int_as_float(thi);                            // no rejected candidate is closer
                for (int e = 0; e < b0; ++e) rest = fminf(rest, bd[e * 32]);
                if (W > 1) for (int e = KS_BND - b1; e < KS_BND; ++e) rest = fminf(rest, bd[e * 32]);
                const float ea = kf_err(a, eA, eC);
                const bool c1 = rest - a > 2.0f * kf_err(rest, eA, eC), c2 = a + ea <= rho1f;
                ok = c1 && c2 && a0 + a1 == c0 && b0 + b1 == cb;
                capf = a + ea * 1.0001f;
                if (!c1) why = 1; else if (!c2) why = 2;
            }
            if (window) { ok = false; why = 5; }      // rho1f < 0 already refuses; the bound capf is what counts
#if KS_W > 1
            sm.ok[lane] = ok ? 1 : 0; sm.cn_up[lane] = (unsigned char)cn_up; sm.cn_dn[lane] = (unsigned char)cn_dn;
#endif
        }
        ks_sync<W>();
#if KS_W > 1
        ok = sm.ok[lane] != 0; cn_up = sm.cn_up[lane]; cn_dn = sm.cn_dn[lane];
#endif
    } else if (nq >= K) {
        capf = U + kf_err(U, eA, eC) * 1.0001f;
    }
    // OUTPUT: sorted positions of the selected neighbours, [j][query position]; KNN_NONE beyond the count
    const bool put = valid && ok && out.nbr != nullptr;
    if (__any_sync(FULL, put)) {
        uint32_t *dst = out.nbr + mypos;
        const int cn = cn_up + cn_dn;
        const int shift = KS_LIST - cn;                   // entry of neighbour j: j below cn_up, else from the top of the column
        const unsigned short *col = sm.v.nb + lane;
        dst += (int64_t)w * out.nbr_pitch;
        for (int j = w; j < K; j += W, dst += (int64_t)W * out.nbr_pitch) {
            const int e = j < cn_up ? j : min(j + shift, KS_LIST - 1);
            const uint32_t pos = j < cn ? (uint32_t)kl_pos(sm.cand, col[e * 32]) : KNN_NONE;
            if (put) *dst = pos;
        }
    }
    if (w == 0) {
        if (valid) {
            capbuf[mypos] = !want ? KF_SKIP : ok ? -fmaxf(capf, 1e-30f) : capf;   // negative = answered, magnitude bounds the k-th d2
            if (out.keep && ok && want) out.tier[mypos] = 1;
        }
        if (out.stats) {
            const unsigned s1 = __reduce_add_sync(FULL, (unsigned)(valid && why == 1)), s2 = __reduce_add_sync(FULL, (unsigned)(valid && why == 2)),
                           s3 = __reduce_add_sync(FULL, (unsigned)(valid && why == 3)), s4 = __reduce_add_sync(FULL, (unsigned)(valid && why == 4)),
                           s5 = __reduce_add_sync(FULL, (unsigned)(valid && why == 5));
            if (lane == 0) {
                atomicAdd(out.stats + 16, 1ull);
                atomicAdd(out.stats + 17, (unsigned long long)(have ? n_lo + n_hi : 32));
                atomicAdd(out.stats + 18, (unsigned long long)scanned);
                atomicAdd(out.stats + 19, (unsigned long long)s1);
                atomicAdd(out.stats + 20, (unsigned long long)s2);
                atomicAdd(out.stats + 21, (unsigned long long)s3);
                atomicAdd(out.stats + 22, (unsigned long long)s4);
                atomicAdd(out.stats + 23, (unsigned long long)s5);
                atomicAdd(out.stats + 24, (unsigned long long)restage);
                const int bk = min((have ? n_lo + n_hi : 32) / 96, 7);          // by staged candidates: where the radius test fails
                atomicAdd(out.stats + 40 + bk, 1ull);
                atomicAdd(out.stats + 32 + bk, (unsigned long long)s2);
            }
        }
    }
}
