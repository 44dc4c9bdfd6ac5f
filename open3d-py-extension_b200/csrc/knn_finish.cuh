// knn_finish.cuh — what follows the neighbour selection of estimate_normals (reference call: PointCloud.py:68-73;
// Open3D EstimateNormals tail: ComputeCovariance + FastEigen3x3 + sign alignment).
//
//   kf_store_normal     covariance -> normal -> sign alignment -> store, with the order-sensitivity guard
//   kf_slow_query       the exact FP64 (d2, index) list search of knn.cuh for ONE query, finished on the spot (tier 3)
//   kf_err              rigorous bound of |d2_fp32 - d2_fp64| for the staged FP32 coordinates of the selection kernels
//   knn_finish_kernel   thread per query: the K neighbour positions the selection kernel (knn_lane.cuh) left in global
//                       memory -> FP64 cumulants -> normal.  It is a separate launch because the selection kernel is
//                       bound by shared memory (few resident warps) while this one is a latency-bound gather
//                       that wants every warp slot of the SM.
//   nbr_fetch_kernel    the recorded neighbour sets as sorted original indices (o3dpp_fetch_neighbors: parity tests compare
//                       the sets of every tier with the oracle's KD-tree)
#pragma once
#include "knn.cuh"
#include <float.h>

#define KF_BIG 1.0e18f
#define KF_HUGE 3.0e36f
#define KF_TCAP 1.0e30f
#define KF_U 5.9604645e-8f /* 2^-24 */
#define KNN_NONE 0xffffffffu
#define KF_SKIP (-3.0e38f)   /* bound-buffer value of a point that is not a query (O3DPP_OPT_QUERY_FIRST): "answered", nothing to finish */

#ifndef KL_GUARD
#define KL_GUARD true   // order-sensitivity guard of the covariance (kf_store_normal); false only for A/B timing
#endif
#ifndef KC_UNR
#define KC_UNR 6        // finish: neighbours whose FP64 coordinates are in flight together
#endif

// covariance -> normal -> sign alignment -> store (Open3D EstimateNormals tail).
// guard = the cumulants were NOT summed in the reference's order (neighbours by increasing (d2, index)).  Open3D's
// one-pass covariance E[ab] - E[a]E[b] cancels |mean|^2 against the variance; when that cancellation times the
// FP64 rounding of a 30-term sum, divided by the eigenvalue separation, could move the direction by more than
// 1e-5 rad, the result depends on the summation order: nothing is stored, false is returned and the caller leaves
// the query to the exact tier, which replays the reference's order.
__device__ __forceinline__ bool kf_store_normal(const Cumul &cu, uint32_t qo, const KnnOut &out, bool guard = false)
{
    double C[6];
    cu.cov(C);
    double nv[3], sep = 0.0;
    fast_eigen3x3(C, nv, guard ? &sep : nullptr);
    if (guard && cu.n >= 3) {
        const double dn = (double)cu.n;
        const double ss = fmax(cu.c[0] * cu.c[0], fmax(cu.c[1] * cu.c[1], cu.c[2] * cu.c[2]));   // (n * largest |mean|)^2
        double mc = C[0];
#pragma unroll
        for (int a = 1; a < 6; ++a) mc = fmax(mc, C[a]);
        if (2e-15 * ss > (1e-5 * sep) * (mc * (dn * dn))) return false;
    }
    double nn = sqrt((nv[0] * nv[0] + nv[1] * nv[1]) + nv[2] * nv[2]);
    double ox = 0, oy = 0, oz = 0;
    if (out.has_old_normals) { ox = out.nx[qo]; oy = out.ny[qo]; oz = out.nz[qo]; }
    if (nn == 0.0) {
        if (out.has_old_normals) { nv[0] = ox; nv[1] = oy; nv[2] = oz; }
        else { nv[0] = 0; nv[1] = 0; nv[2] = 1; }
    }
    if (out.has_old_normals && ((nv[0] * ox + nv[1] * oy) + nv[2] * oz) < 0.0) { nv[0] = -nv[0]; nv[1] = -nv[1]; nv[2] = -nv[2]; }
    out.nx[qo] = nv[0]; out.ny[qo] = nv[1]; out.nz[qo] = nv[2];
    return true;
}

// exact answer of one query by the FP64 range-table search (the certified path's fallback), finished
// on the spot: every lane replays the (d2, index)-ordered cumulants, lane 0 stores the normal.
// cap = a known upper bound of the k-th d2 (+inf if none)
template <int MODE>
__device__ __noinline__ void kf_slow_query(const GridDesc &g, int64_t leaf_id, int i, int K, double r2, double cap,
                                           const KnnOut &out, float *kth_out = nullptr)
{
    const int lane = lane_id();
    const int64_t mypos = leaf_id * 32 + lane;
    const bool valid = mypos < g.n;
    double px = 0, py = 0, pz = 0;
    if (valid) { px = g.sx[mypos]; py = g.sy[mypos]; pz = g.sz[mypos]; }
    Search<MODE, 1> st(g);
    st.lane = lane; st.K = K; st.r2 = r2; st.home = (int)leaf_id; st.cap = cap;
    st.st_visit = st.st_leaf = st.st_ins = st.st_cand = 0;
    st.qx = shfl_d(px, i); st.qy = shfl_d(py, i); st.qz = shfl_d(pz, i);
    double d = valid ? dist2_rn(st.qx, st.qy, st.qz, px, py, pz) : pos_inf();
    int p = valid ? (int)mypos : -1;
    bitonic_sort32<1>(d, p, lane, g.sidx);
    st.L.d2[0] = d; st.L.pos[0] = p;
    st.refresh_worst();
    st.walk();
    bool ok = lane < K && st.L.pos[0] >= 0 && (MODE != 2 || st.L.d2[0] < r2);
    int c = __popc(__ballot_sync(FULL, ok));
    if (kth_out) {      // the exact k-th d2 (rounded up), negated: "answered" in the lane path's bound buffer
        double kd = shfl_d(st.L.d2[0], c > 0 ? c - 1 : 0);
        if (lane == 0) *kth_out = -fmaxf(c > 0 ? __double2float_ru(kd) : 0.f, 1e-30f);
    }
    if (out.keep) {     // neighbour-set record (o3dpp_fetch_neighbors): sorted positions, KNN_NONE beyond the count
        const int64_t qpos = leaf_id * 32 + i;
        if (lane < K) out.nbr[(int64_t)lane * out.nbr_pitch + qpos] = lane < c ? (uint32_t)st.L.pos[0] : KNN_NONE;
        if (lane == 0) out.tier[qpos] = 3;
    }
    if (!out.nx) return;
    // the reference's summation order (neighbours by increasing (d2, index) = lane order): every lane fetches ITS neighbour once,
    // the replay broadcasts them in turn (one gather instead of c dependent round trips: a tier-3 query is a single warp's
    // latency chain, 30 serial loads were a third of it)
    double vx = 0, vy = 0, vz = 0;
    if (lane < c) { const double4 r = ld_rec(g.sp4 + st.L.pos[0]); vx = r.x; vy = r.y; vz = r.z; }
    Cumul cu;
    cu.clear();
    for (int j = 0; j < c; ++j) cu.add(shfl_d(vx, j), shfl_d(vy, j), shfl_d(vz, j));
    if (lane == 0) kf_store_normal(cu, g.sidx[leaf_id * 32 + i], out);
    __syncwarp();
}

__device__ __forceinline__ float kf_err(float t, float eA, float eC)
{
    // |d2_fp32 - d2_fp64| <= u (14 d L + 3 d^2) + 49 u^2 L^2 ; eA = 16 u L, eC = 2^-42 L^2, 1 % slack
    return 1.01f * (eA * sqrtf(t) + 4.0f * KF_U * t + eC);
}

// ---- finish: the neighbour positions left by the selection kernel -> FP64 cumulants -> normal -----------------------
// capbuf[q] < 0: answered by the selection kernel, |capbuf| bounds the k-th d2.  A query whose covariance is
// order-sensitive (kf_store_normal's guard) is handed back (capbuf > 0) to the exact tier.
//
// Thread per query, the K gathers in batches of KC_UNR.  Measured at 10 M and not kept: a lane PAIR per 32-byte record (16 + 16
// bytes in one instruction, to halve the L1 wavefronts) 2.16 ms against 1.58 (the kernel is latency bound, not wavefront bound);
// running it on a second stream beside tiers 2 and 3 (it fills the register file: the kernels serialise anyway).
// 64 registers (KC_MINB 8) against 88: 1.30 against 1.58 ms.  Also measured without gain: prefetch.global.L1 of all K records before the
// loop (2.04 ms), FMAs in the cumulants (1.31), reciprocal multiplications for the ~20 divisions by a common denominator in the
// covariance and the eigen-solve (1.40 ms, and the 1e-7 agreement with the oracle on well-conditioned points is lost).
#ifndef KC_MINB
#define KC_MINB 8       // resident 128-thread blocks per SM the register budget is set for
#endif
template <int MODE>
__global__ void __launch_bounds__(128, KC_MINB)
knn_finish_kernel(const GridDesc g, int K, double radius, KnnOut out, float *__restrict__ capbuf)
{
    const int64_t q = (int64_t)blockIdx.x * 128 + threadIdx.x;
    if (q >= g.n) return;
    const float cap = capbuf[q];
    if (!(cap < 0.0f) || cap <= KF_SKIP) return;
    const double r2 = __dmul_rn(radius, radius);
    double px = 0, py = 0, pz = 0;
    if (MODE == 2) { const double4 me = g.sp4[q]; px = me.x; py = me.y; pz = me.z; }
    const uint32_t *nb = out.nbr + q;
    Cumul cu;
    cu.clear();
    for (int j0 = 0; j0 < K; j0 += KC_UNR) {
        uint32_t p[KC_UNR];
        double4 v[KC_UNR];
#pragma unroll
        for (int u = 0; u < KC_UNR; ++u) p[u] = j0 + u < K ? nb[(int64_t)(j0 + u) * out.nbr_pitch] : KNN_NONE;
#pragma unroll
        for (int u = 0; u < KC_UNR; ++u) v[u] = ld_rec(g.sp4 + (p[u] != KNN_NONE ? (int64_t)p[u] : q));     // unconditional: all in flight
#pragma unroll
        for (int u = 0; u < KC_UNR; ++u)
            if (p[u] != KNN_NONE && (MODE != 2 || dist2_rn(px, py, pz, v[u].x, v[u].y, v[u].z) < r2)) cu.add(v[u].x, v[u].y, v[u].z);
    }
    if (!kf_store_normal(cu, g.sidx[q], out, KL_GUARD)) capbuf[q] = -cap;
}

// ---- recorded neighbour sets -> sorted original indices ------------------------------------------------------------
__global__ void __launch_bounds__(256)
inverse_perm_kernel(const uint32_t *__restrict__ sidx, int64_t n, uint32_t *__restrict__ inv)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) inv[sidx[i]] = (uint32_t)i;
}

// thread per requested query (original index; queries == nullptr: all points in order): the K recorded positions ->
// original indices, hybrid radius filter in FP64 (the selection tier leaves it to the finish), ascending, -1 padded
__global__ void __launch_bounds__(128)
nbr_fetch_kernel(const GridDesc g, const uint32_t *__restrict__ nbr, int64_t pitch, const unsigned char *__restrict__ tier,
                 const uint32_t *__restrict__ inv, const int64_t *__restrict__ queries, int64_t m, int K, int mode, double radius,
                 int32_t *__restrict__ idx, int32_t *__restrict__ cnt, int8_t *__restrict__ tier_out)
{
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= m) return;
    const int64_t o = queries ? queries[t] : t;
    int32_t *row = idx + t * K;
    if (o < 0 || o >= g.n) {
        for (int j = 0; j < K; ++j) row[j] = -1;
        if (cnt) cnt[t] = -1;
        if (tier_out) tier_out[t] = -1;
        return;
    }
    const uint32_t q = inv[o];
    const double4 me = g.sp4[q];
    const double r2 = __dmul_rn(radius, radius);
    int c = 0;
    for (int j = 0; j < K; ++j) {
        const uint32_t p = nbr[(int64_t)j * pitch + q];
        if (p == KNN_NONE) continue;
        if (mode == O3DPP_SEARCH_HYBRID) {
            const double4 v = g.sp4[p];
            if (!(dist2_rn(me.x, me.y, me.z, v.x, v.y, v.z) < r2)) continue;
        }
        const int32_t oi = (int32_t)g.sidx[p];
        int a = c;
        while (a > 0 && row[a - 1] > oi) { row[a] = row[a - 1]; --a; }
        row[a] = oi;
        ++c;
    }
    for (int j = c; j < K; ++j) row[j] = -1;
    if (cnt) cnt[t] = c;
    if (tier_out) tier_out[t] = tier ? (int8_t)tier[q] : 0;
}
