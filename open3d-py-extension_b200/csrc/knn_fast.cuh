// knn_fast.cuh — the estimate_normals fast path: certified FP32 selection over staged candidate
// rows, exact FP64 covariance + eigen-solve.  (Reference call: PointCloud.py:68-73.)
//
// One warp owns one cell group (32 Hilbert-consecutive points = 32 queries) and works in three steps:
//
//  1. STAGE (once per warp).  The range tables are walked for the whole group: every cell group whose
//     box lies within rho of the home box is scanned, and its points that fall inside the home box
//     grown by rho are appended to shared memory as FP32 coordinates relative to the home-box centre
//     (recentred in FP64 first), 32 to a row, with an FP32 box per row.  Every point within rho of any
//     query of the home box is therefore staged.
//  2. SELECT (one query at a time, 32 lanes on its candidates).  Lane r holds the box distance of row
//     r; rows whose box distance exceeds the current threshold t are never touched.  FP32 d2 of the
//     active rows live in registers.  The k-th smallest d2 is bracketed by counting: `count(d2 <= t)`
//     is a predicated add per register and one REDUX, first over the active rows until
//     k < count <= 96, then over the <= 96 survivors (three per lane) until count == k.  The result
//     is a SET; no sorted list is maintained.
//  3. CERTIFY.  With a = k-th and b = (k+1)-th FP32 d2, the set equals the exact FP64 (d2, index) set
//     if b - a > 2E, E a rigorous bound of |d2_fp32 - d2_fp64| (recentring + rounding), and it is
//     complete if a + E <= rho^2.  A query failing either test (near ties, duplicates, outliers) is
//     answered by the exact FP64 range-table search of knn.cuh, seeded with a + E as its bound, so
//     the result is always the exact (d2, index) neighbour set.
//
// After the 32 queries every lane finishes ONE query: FP64 single-pass cumulants over the selected
// neighbours (Open3D ComputeCovariance), closed-form eigenvector, sign alignment, store.
#pragma once
#include "knn.cuh"
#include <float.h>

#define KF_ROWS 24    // staged rows of 32 candidates (lane r keeps the box of row r)
#define KF_CAP (KF_ROWS * 32)
#define KF_AMAX 16    // rows one query may activate (their d2 live in registers)
#define KF_SURV 96    // survivors of phase A, three per lane
#define KF_WARPS 4
#define KF_BIG 1.0e18f
#define KF_HUGE 3.0e36f
#define KF_TCAP 1.0e30f
#define KF_U 5.9604645e-8f /* 2^-24 */

struct KfWarpMem {
    float x[KF_CAP], y[KF_CAP], z[KF_CAP];
    int pos[KF_CAP];                     // sorted position of every staged candidate
    unsigned long long surv[KF_SURV];    // (fp32 d2 bits << 32) | staged slot
    int nb[32 * 33];                     // selected neighbours (sorted positions) of the 32 queries
    int cnt[32];
};

__device__ __forceinline__ int kf_ord(float f)
{
    int b = __float_as_int(f);
    return b ^ ((b >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float kf_unord(int o) { return __int_as_float(o ^ ((o >> 31) & 0x7fffffff)); }

struct KfStage {
    const GridDesc &g;
    KfWarpMem &sm;
    int lane;
    double blo[3], bhi[3];      // home box
    double cc[3];               // its centre
    double rho, rho2;
    int n;                      // staged candidates
    bool ball;                  // home box is a point: keep the points within rho of it (else: box grown by rho)
    bool overflow;
    unsigned scanned;
    __device__ __forceinline__ KfStage(const GridDesc &g_, KfWarpMem &sm_) : g(g_), sm(sm_) {}

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
            if (ball) in = (lx * lx + ly * ly) + lz * lz <= rho2;
            else in = x >= blo[0] - rho && x <= bhi[0] + rho && y >= blo[1] - rho && y <= bhi[1] + rho &&
                      z >= blo[2] - rho && z <= bhi[2] + rho;
        }
        unsigned m = __ballot_sync(FULL, in);
        int at = n + __popc(m & ((1u << lane) - 1u));
        if (in && at < KF_CAP) { sm.x[at] = (float)lx; sm.y[at] = (float)ly; sm.z[at] = (float)lz; sm.pos[at] = (int)p; }
        n += __popc(m);
        if (n > KF_CAP) overflow = true;
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
            if constexpr (LVL == 0) scan_group(parent * 32 + b);
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
    if (!out.nx) return;
    Cumul cu;
    cu.clear();
    for (int j = 0; j < c; ++j) {
        int q = __shfl_sync(FULL, st.L.pos[0], j);
        cu.add(g.sx[q], g.sy[q], g.sz[q]);
    }
    if (lane == 0) kf_store_normal(cu, g.sidx[leaf_id * 32 + i], out);
    __syncwarp();
}

__device__ __forceinline__ float kf_err(float t, float eA, float eC)
{
    // |d2_fp32 - d2_fp64| <= u (14 d L + 3 d^2) + 49 u^2 L^2 ; eA = 16 u L, eC = 2^-42 L^2, 1 % slack
    return 1.01f * (eA * sqrtf(t) + 4.0f * KF_U * t + eC);
}

struct KfStats { unsigned slow, t2, ra, rc, fw, fa, fc, fgap, frho, act, rows, scanned; };

// what every lane knows about the staged rows
struct KfRows {
    float bx0, by0, bz0, bx1, by1, bz1;   // FP32 box of row `lane`
    int nrows, ncand;
    float eA, eC;                          // error model of the staged FP32 coordinates
    float rho1f;                           // everything within sqrt(rho1f) of the home box is staged
    float tcap;                            // largest useful threshold (hybrid: radius^2 + slack)
};

// rows are in shared memory (ncand candidates): pad the last row, give lane r the FP32 box of row r,
// derive the error model.  rho1sq < 0 = nothing is guaranteed about what was left out.
template <int MODE>
__device__ __forceinline__ bool kf_rows_finish(KfWarpMem &sm, int lane, bool ok, int ncand, double rho1sq, double r2d, KfRows &R)
{
    R.ncand = ok ? ncand : 0;
    R.nrows = (R.ncand + 31) >> 5;
    R.rho1f = (ok && rho1sq > 0.0) ? __double2float_rd(rho1sq * (1.0 - 1e-6)) : -1.0f;
    if (ok) {   // pad the last row
        int padat = R.ncand + lane;
        if (padat < R.nrows * 32) { sm.x[padat] = KF_BIG; sm.y[padat] = KF_BIG; sm.z[padat] = KF_BIG; sm.pos[padat] = -1; }
    }
    __syncwarp();
    R.bx0 = R.by0 = R.bz0 = KF_BIG; R.bx1 = R.by1 = R.bz1 = -KF_BIG;
    float lm = 0.f;
    if (ok) {
        for (int r = 0; r < R.nrows; ++r) {
            int s = r * 32 + lane;
            bool v = s < R.ncand;
            int ox = kf_ord(sm.x[s]), oy = kf_ord(sm.y[s]), oz = kf_ord(sm.z[s]);
            int ox0 = __reduce_min_sync(FULL, v ? ox : 0x7fffffff), ox1 = __reduce_max_sync(FULL, v ? ox : (int)0x80000000);
            int oy0 = __reduce_min_sync(FULL, v ? oy : 0x7fffffff), oy1 = __reduce_max_sync(FULL, v ? oy : (int)0x80000000);
            int oz0 = __reduce_min_sync(FULL, v ? oz : 0x7fffffff), oz1 = __reduce_max_sync(FULL, v ? oz : (int)0x80000000);
            if (lane == r) { R.bx0 = kf_unord(ox0); R.bx1 = kf_unord(ox1); R.by0 = kf_unord(oy0); R.by1 = kf_unord(oy1); R.bz0 = kf_unord(oz0); R.bz1 = kf_unord(oz1); }
        }
        if (lane < R.nrows)
            lm = fmaxf(fmaxf(fmaxf(fabsf(R.bx0), fabsf(R.bx1)), fmaxf(fabsf(R.by0), fabsf(R.by1))), fmaxf(fabsf(R.bz0), fabsf(R.bz1)));
        lm = __uint_as_float(__reduce_max_sync(FULL, __float_as_uint(lm)));
        if (!(lm < 1e15f)) ok = false;
    }
    R.eA = 16.0f * KF_U * lm; R.eC = 2.3e-13f * lm * lm;
    R.tcap = KF_TCAP;
    if (MODE == 2) {
        float r2f = __double2float_ru(r2d);
        R.tcap = r2f * (1.0f + 1e-6f) + 2.0f * kf_err(r2f, R.eA, R.eC);
        if (!(R.tcap < KF_TCAP)) ok = false;
    }
    return ok;
}

// Stage everything within `rho` (shrunk on overflow) of sg's home box.  false = cannot be staged.
template <int MODE>
__device__ __forceinline__ bool kf_stage(KfStage &sg, KfWarpMem &sm, double rho, double r2d, KfRows &R)
{
    bool ok = true;
    for (int attempt = 0;; ++attempt) {
        sg.n = 0; sg.overflow = false; sg.rho = rho; sg.rho2 = rho * rho * (1.0 + 1e-12);
        sg.walk();
        if (!sg.overflow) break;
        if (attempt >= (sg.ball ? 2 : 8)) { ok = false; break; }
        rho *= 0.7;
    }
    ok = kf_rows_finish<MODE>(sm, sg.lane, ok, sg.n, rho * rho, r2d, R);
    return ok && R.rho1f > 0.0f;
}

// Select the K nearest staged candidates of the query at local (qx,qy,qz).  true = certified exact:
// nbrow[0..cnt) holds their sorted positions.  cap (out) = upper bound of the k-th d2 when known.
template <int MODE>
__device__ __forceinline__ bool kf_select(KfWarpMem &sm, const KfRows &R, float qx, float qy, float qz, int K, float &tg,
                                          int *nbrow, int *cntp, double &cap, KfStats &st)
{
    const int lane = lane_id();
    const unsigned lt = (1u << lane) - 1u;
    // box distance of row `lane`, same expression as the point distance so that lbr <= d2 of every member
    float lbr = KF_HUGE;
    if (lane < R.nrows) {
        float gx = fmaxf(fmaxf(R.bx0 - qx, qx - R.bx1), 0.f), gy = fmaxf(fmaxf(R.by0 - qy, qy - R.by1), 0.f), gz = fmaxf(fmaxf(R.bz0 - qz, qz - R.bz1), 0.f);
        lbr = fmaf(gz, gz, fmaf(gy, gy, gx * gx));
    }
    float d2[KF_AMAX];
    int rowof[KF_AMAX];
#pragma unroll
    for (int a = 0; a < KF_AMAX; ++a) { d2[a] = KF_HUGE; rowof[a] = 0; }
    unsigned act_done = 0;
    int nact = 0;
    // phase A: t with K < count(d2 <= t) <= KF_SURV
    float t = fminf(tg, R.tcap), lo = 0.f, hi = FLT_MAX;
    int c = 0;
    bool all = false, takeall = false;   // all: the answer is every survivor; takeall: exactly K are staged
    for (int it = 0;; ++it) {
        unsigned act = __ballot_sync(FULL, lbr <= t);
        unsigned newm = act & ~act_done;
        if (__popc(act | act_done) > KF_AMAX) { st.fa++; return false; }
        act_done |= act;
#pragma unroll
        for (int a = 0; a < KF_AMAX; ++a) {
            if (newm != 0u && a == nact) {
                int r = __ffs(newm) - 1;
                newm &= newm - 1;
                int s = r * 32 + lane;
                float dx = sm.x[s] - qx, dy = sm.y[s] - qy, dz = sm.z[s] - qz;
                d2[a] = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                rowof[a] = r;
                nact++;
            }
        }
        int cl = 0;
#pragma unroll
        for (int a = 0; a < KF_AMAX; ++a) cl += d2[a] <= t ? 1 : 0;
        c = __reduce_add_sync(FULL, cl);
        st.ra++;
        if (c > K && c <= KF_SURV) break;
        if (it >= 14) { st.fa++; return false; }
        if (c <= K) {
            if (c >= R.ncand) {             // every staged candidate is already inside t
                if (c == K || (MODE == 2 && t >= R.tcap)) { all = true; takeall = c == K && !(MODE == 2 && t >= R.tcap); break; }
                st.fa++;
                return false;
            }
            if (t >= R.tcap) {
                if (MODE == 2) { all = true; break; }
                st.fa++;
                return false;
            }
            lo = t;
            float tn = t * fminf(4.0f, 56.0f / (float)(c > 0 ? c : 1));
            if (tn >= hi) tn = 0.5f * (lo + hi);
            if (!(tn > lo)) { st.fa++; return false; }
            t = fminf(tn, R.tcap);
        } else {
            hi = t;
            float tn = t * fmaxf(0.1f, 56.0f / (float)c);
            if (tn <= lo) tn = 0.5f * (lo + hi);
            if (!(tn < hi)) { st.fa++; return false; }
            t = tn;
        }
    }
    st.act += nact;
    // survivors, three per lane
    int nb_ = 0;
#pragma unroll
    for (int a = 0; a < KF_AMAX; ++a) {
        if (a < nact) {
            bool pass = d2[a] <= t;
            unsigned m = __ballot_sync(FULL, pass);
            if (pass)
                sm.surv[nb_ + __popc(m & lt)] = ((unsigned long long)__float_as_uint(d2[a]) << 32) | (unsigned)(rowof[a] * 32 + lane);
            nb_ += __popc(m);
        }
    }
    __syncwarp();
    unsigned long long e0 = lane < c ? sm.surv[lane] : ~0ull;
    unsigned long long e1 = lane + 32 < c ? sm.surv[lane + 32] : ~0ull;
    unsigned long long e2 = lane + 64 < c ? sm.surv[lane + 64] : ~0ull;
    __syncwarp();
    const unsigned k0 = (unsigned)(e0 >> 32), k1 = (unsigned)(e1 >> 32), k2 = (unsigned)(e2 >> 32);
    unsigned akey = 0u, bkey = 0xffffffffu;
    bool okc = all;
    if (!all) {
        // phase C: pivot with count(key <= pivot) == K ; invariant count(<=lb) < K < count(<=ub)
        unsigned lb = 0u, ub = __float_as_uint(t);
        int cl_ = 0, ch_ = c;
        unsigned piv = __float_as_uint(t * (((float)K + 0.5f) / (float)c));
        if (piv >= ub) piv = ub - 1u;
        for (int it = 0; it < 32; ++it) {
            st.rc++;
            bool s0 = k0 <= piv, s1 = k1 <= piv, s2 = k2 <= piv;
            int cnt = __reduce_add_sync(FULL, (int)s0 + (int)s1 + (int)s2);
            unsigned am = max(max(s0 ? k0 : 0u, s1 ? k1 : 0u), s2 ? k2 : 0u);
            unsigned bm = min(min(s0 ? 0xffffffffu : k0, s1 ? 0xffffffffu : k1), s2 ? 0xffffffffu : k2);
            akey = __reduce_max_sync(FULL, am);
            bkey = __reduce_min_sync(FULL, bm);
            if (cnt == K) { okc = true; break; }
            if (cnt > K) { ub = akey; ch_ = cnt; }
            else { lb = bkey - 1u; cl_ = cnt; }
            if (ub <= lb + 1u) break;     // the remaining candidates tie in FP32
            if (cnt > K && cnt - K <= 2) piv = ub - 1u;
            else if (cnt < K && K - cnt <= 2) piv = bkey;
            else {
                float lf = __uint_as_float(lb), uf = __uint_as_float(ub);
                piv = __float_as_uint(lf + (uf - lf) * ((float)(K - cl_) / (float)(ch_ - cl_)));
            }
            if (piv < lb + 1u) piv = lb + 1u;
            if (piv > ub - 1u) piv = ub - 1u;
        }
    }
    if (!okc) { st.fc++; return false; }
    bool cert;
    float af;
    if (all) {
        af = R.tcap;
        if (takeall) {
            unsigned mk = max(max(lane < c ? k0 : 0u, lane + 32 < c ? k1 : 0u), lane + 64 < c ? k2 : 0u);
            af = __uint_as_float(__reduce_max_sync(FULL, mk));
            cap = (double)af + (double)kf_err(af, R.eA, R.eC) * 1.0001;
        }
        cert = af + kf_err(af, R.eA, R.eC) <= R.rho1f;
        if (!cert) st.frho++;
    } else {
        af = __uint_as_float(akey);
        float bf = __uint_as_float(bkey);
        float ea = kf_err(af, R.eA, R.eC);
        bool c1 = bf - af > 2.0f * kf_err(bf, R.eA, R.eC), c2 = af + ea <= R.rho1f;
        cert = c1 && c2;
        if (!c1) st.fgap++;
        else if (!c2) st.frho++;
        cap = (double)af + (double)ea * 1.0001;
    }
    if (!cert) return false;
    bool s0 = all ? (lane < c) : (k0 <= akey), s1 = all ? (lane + 32 < c) : (k1 <= akey), s2 = all ? (lane + 64 < c) : (k2 <= akey);
    unsigned m0 = __ballot_sync(FULL, s0), m1 = __ballot_sync(FULL, s1), m2 = __ballot_sync(FULL, s2);
    if (s0) nbrow[__popc(m0 & lt)] = sm.pos[(unsigned)e0 & 0xffffu];
    if (s1) nbrow[__popc(m0) + __popc(m1 & lt)] = sm.pos[(unsigned)e1 & 0xffffu];
    if (s2) nbrow[__popc(m0) + __popc(m1) + __popc(m2 & lt)] = sm.pos[(unsigned)e2 & 0xffffu];
    if (lane == 0) *cntp = __popc(m0) + __popc(m1) + __popc(m2);
    tg = fmaxf(af * 1.6f, 1e-30f);
    return true;
}

// Bound step of the single-query search: descend the range tables towards the entry whose farthest
// corner is nearest to q (every entry with >= K points bounds the k-th distance by that corner), then
// stage the three best cell groups of the last node unfiltered as rows 0..2.  Returns the staged count.
__device__ __forceinline__ int kf_bound_rows(const GridDesc &g, KfWarpMem &sm, int lane, int K, double qx, double qy, double qz,
                                             int64_t qpos, float *ub2_best)
{
    int64_t node = 0;
    int64_t span = 32;                       // points per entry of the current table
    for (int l = 1; l < g.levels; ++l) span *= 32;
    unsigned pick[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu};
    for (int lvl = g.levels - 1; lvl >= 0; --lvl, span >>= 5) {
        int64_t c = node * 32 + lane;
        float ub = FLT_MAX;
        if (c < g.cnt[lvl]) {
            int64_t have = g.n - c * span;
            if (have > span) have = span;
            if (have >= K) {
                const double *b = g.box + g.off[lvl];
                int64_t p = g.pitch[lvl];
                double fx = fmax(qx - b[c], b[3 * p + c] - qx), fy = fmax(qy - b[p + c], b[4 * p + c] - qy), fz = fmax(qz - b[2 * p + c], b[5 * p + c] - qz);
                ub = __double2float_ru((fx * fx + fy * fy) + fz * fz);
            }
        }
        unsigned ubits = __float_as_uint(ub);
        unsigned best = __reduce_min_sync(FULL, ubits);
        if (best == __float_as_uint(FLT_MAX)) { *ub2_best = 0.f; break; }     // nothing holds K points (tiny cloud)
        int bl = __ffs(__ballot_sync(FULL, ubits == best)) - 1;
        if (lvl > 0) { node = node * 32 + bl; continue; }
        *ub2_best = __uint_as_float(best);
        pick[0] = (unsigned)(node * 32 + bl);
        for (int j = 1; j < 3; ++j) {
            if (lane == bl) ubits = 0xffffffffu;
            best = __reduce_min_sync(FULL, ubits);
            if (best >= __float_as_uint(FLT_MAX)) break;
            bl = __ffs(__ballot_sync(FULL, ubits == best)) - 1;
            pick[j] = (unsigned)(node * 32 + bl);
        }
    }
    int rows = 0;
    for (int j = 0; j < 3; ++j) {
        if (pick[j] == 0xffffffffu) break;
        int64_t p = (int64_t)pick[j] * 32 + lane;
        int s = rows * 32 + lane;
        if (p < g.n) { sm.x[s] = (float)(g.sx[p] - qx); sm.y[s] = (float)(g.sy[p] - qy); sm.z[s] = (float)(g.sz[p] - qz); sm.pos[s] = (int)p; }
        else { sm.x[s] = KF_BIG; sm.y[s] = KF_BIG; sm.z[s] = KF_BIG; sm.pos[s] = -1; }
        rows++;
    }
    // plus the 96 points around q along the curve (skipping the groups staged above: no point twice)
    int64_t w0 = qpos - 48;
    if (w0 + 96 > g.n) w0 = g.n - 96;
    if (w0 < 0) w0 = 0;
    for (int j = 0; j < 3; ++j) {
        int64_t p = w0 + j * 32 + lane;
        int s = rows * 32 + lane;
        unsigned lf = (unsigned)(p >> 5);
        bool v = p < g.n && lf != pick[0] && lf != pick[1] && lf != pick[2];
        if (v) { sm.x[s] = (float)(g.sx[p] - qx); sm.y[s] = (float)(g.sy[p] - qy); sm.z[s] = (float)(g.sz[p] - qz); sm.pos[s] = (int)p; }
        else { sm.x[s] = KF_BIG; sm.y[s] = KF_BIG; sm.z[s] = KF_BIG; sm.pos[s] = -1; }
        rows++;
    }
    __syncwarp();
    return rows * 32;
}

template <int MODE>
__global__ void __launch_bounds__(KF_WARPS * 32)
knn_fast_kernel(const GridDesc g, int K, double radius, float rho_scale, KnnOut out, int dbg_skip)
{
    extern __shared__ __align__(16) unsigned char kf_smem[];
    const int lane = lane_id(), w = threadIdx.x >> 5;
    const int64_t leaf_id = (int64_t)blockIdx.x * KF_WARPS + w;
    if (leaf_id >= g.cnt[0]) return;
    KfWarpMem &sm = ((KfWarpMem *)kf_smem)[w];
    const int64_t base = leaf_id * 32;
    const int nq = (int)(g.n - base < 32 ? g.n - base : 32);
    const double r2d = __dmul_rn(radius, radius);
    KfStats st;
    memset(&st, 0, sizeof st);

    // own query, FP64
    const int64_t mypos = base + lane;
    double px = 0, py = 0, pz = 0;
    if (mypos < g.n) { px = g.sx[mypos]; py = g.sy[mypos]; pz = g.sz[mypos]; }

    // ---- tier 1: the whole group against one staged set ----------------------------------------
    KfStage sg(g, sm);
    sg.lane = lane; sg.scanned = 0; sg.ball = false;
    {
        const double *b0 = g.box + g.off[0];
        int64_t p0 = g.pitch[0];
        sg.blo[0] = b0[leaf_id]; sg.blo[1] = b0[p0 + leaf_id]; sg.blo[2] = b0[2 * p0 + leaf_id];
        sg.bhi[0] = b0[3 * p0 + leaf_id]; sg.bhi[1] = b0[4 * p0 + leaf_id]; sg.bhi[2] = b0[5 * p0 + leaf_id];
    }
    sg.cc[0] = 0.5 * (sg.blo[0] + sg.bhi[0]); sg.cc[1] = 0.5 * (sg.blo[1] + sg.bhi[1]); sg.cc[2] = 0.5 * (sg.blo[2] + sg.bhi[2]);
    double rho_g;
    {
        double ex = sg.bhi[0] - sg.blo[0], ey = sg.bhi[1] - sg.blo[1], ez = sg.bhi[2] - sg.blo[2];
        rho_g = sqrt((ex * ex + ey * ey) + ez * ez) * (double)rho_scale;
        if (MODE == 2 && radius * 1.001 < rho_g) rho_g = radius * 1.001;
    }
    KfRows R;
    const bool group_ok = kf_stage<MODE>(sg, sm, rho_g, r2d, R);
    st.rows = R.nrows;
    // own query in local FP32 (bit-identical to its staged copy)
    const float mqx = (float)(px - sg.cc[0]), mqy = (float)(py - sg.cc[1]), mqz = (float)(pz - sg.cc[2]);
    float tg = 0.25f * R.rho1f;
    unsigned failed = 0;       // queries tier 1 could not certify
    double mycap = pos_inf();  // lane i: upper bound of query i's k-th d2
    for (int i = 0; i < nq; ++i) {
        bool done = false;
        double cap = pos_inf();
        if (group_ok) {
            const float qx = __shfl_sync(FULL, mqx, i), qy = __shfl_sync(FULL, mqy, i), qz = __shfl_sync(FULL, mqz, i);
            done = kf_select<MODE>(sm, R, qx, qy, qz, K, tg, sm.nb + i * 33, sm.cnt + i, cap, st);
        } else st.fw++;
        if (!done) { failed |= 1u << i; if (lane == i) mycap = cap; }
    }
    __syncwarp();

    // ---- tier 2: one failed query at a time against its own staged ball ---------------------------
    if (dbg_skip & 1) { for (int i = 0; i < 32; ++i) if ((failed >> i & 1) && lane == 0) sm.cnt[i] = 0; failed = 0; }   // timing experiments only
    while (failed) {
        const int i = __ffs(failed) - 1;
        failed &= failed - 1;
        st.t2++;
        const double qx = shfl_d(px, i), qy = shfl_d(py, i), qz = shfl_d(pz, i);
        const double capi = shfl_d(mycap, i);
        sg.ball = true;
        sg.blo[0] = sg.bhi[0] = sg.cc[0] = qx; sg.blo[1] = sg.bhi[1] = sg.cc[1] = qy; sg.blo[2] = sg.bhi[2] = sg.cc[2] = qz;
        double capq = capi;
        if (!(dbg_skip & 4)) {   // bound step: k-th d2 among the three cell groups nearest to q (by farthest corner)
            float ubb = 0.f;
            int nb3 = kf_bound_rows(g, sm, lane, K, qx, qy, qz, base + i, &ubb);
            if (nb3 > 0) {
                KfRows Rb;
                kf_rows_finish<MODE>(sm, lane, true, nb3, -1.0, r2d, Rb);
                float tb = ubb > 0.f ? ubb : 1.0f;
                double capb = pos_inf();
                int dummy_cnt;
                (void)kf_select<MODE>(sm, Rb, 0.f, 0.f, 0.f, K, tb, sm.nb + i * 33, &dummy_cnt, capb, st);
                if (capb < capq) capq = capb;
            }
            __syncwarp();
        }
        double rho_q = capq < 1e300 ? sqrt(capq) * 1.01 : 2.0 * rho_g;
        if (MODE == 2 && radius * 1.001 < rho_q) rho_q = radius * 1.001;
        bool done = false;
        if (rho_q > 0.0 && !(dbg_skip & 2)) {
            for (int attempt = 0; attempt < 3 && !done; ++attempt) {
                KfRows Rq;
                if (!kf_stage<MODE>(sg, sm, rho_q, r2d, Rq)) break;
                float t2 = capq < 1e300 ? (float)capq * fminf(1.0f, 56.0f / (float)(Rq.ncand > 0 ? Rq.ncand : 1)) : 0.25f * Rq.rho1f;
                double cap2 = pos_inf();
                done = kf_select<MODE>(sm, Rq, 0.f, 0.f, 0.f, K, t2, sm.nb + i * 33, sm.cnt + i, cap2, st);
                if ((dbg_skip & 8) && lane == 0 && leaf_id % 499 == 0)
                    printf("t2 leaf %lld q %d att %d capi %.4g capq %.4g rho_q %.4g -> rho1f %.4g ncand %d rows %d done %d cap2 %.4g scanned_total %u\n",
                           (long long)leaf_id, i, attempt, capi, capq, rho_q, (double)Rq.rho1f, Rq.ncand, Rq.nrows, (int)done, cap2, sg.scanned);
                if (capq < 1e300) break;                   // the ball of a valid bound holds the answer: growing cannot help
                if (MODE == 2 && rho_q >= radius) break;   // the radius ball was staged in full
                rho_q *= 2.0;
            }
        }
        if (!done) {
            kf_slow_query<MODE>(g, leaf_id, i, K, r2d, capq, out);   // tier 3 finishes the query itself
            if (lane == 0) sm.cnt[i] = -1;
            st.slow++;
        }
        __syncwarp();
    }
    st.scanned = sg.scanned;
    if (out.stats && lane == 0) {
        atomicAdd(out.stats + 4, (unsigned long long)st.slow);
        atomicAdd(out.stats + 5, (unsigned long long)st.rows);
        atomicAdd(out.stats + 6, (unsigned long long)st.ra);
        atomicAdd(out.stats + 7, (unsigned long long)st.rc);
        atomicAdd(out.stats + 8, (unsigned long long)st.fw);
        atomicAdd(out.stats + 9, (unsigned long long)st.fa);
        atomicAdd(out.stats + 10, (unsigned long long)st.fc);
        atomicAdd(out.stats + 11, (unsigned long long)st.fgap);
        atomicAdd(out.stats + 12, (unsigned long long)st.frho);
        atomicAdd(out.stats + 13, (unsigned long long)st.scanned);
        atomicAdd(out.stats + 14, (unsigned long long)st.act);
        atomicAdd(out.stats + 15, (unsigned long long)st.t2);
    }
    // ---- every lane finishes one query ------------------------------------------------------------
    if (mypos >= g.n || !out.nx) return;
    const int cn = sm.cnt[lane];
    if (cn < 0) return;   // answered by the exact search
    Cumul cu;
    cu.clear();
    for (int j = 0; j < cn; ++j) {
        int p = sm.nb[lane * 33 + j];
        double x = g.sx[p], y = g.sy[p], z = g.sz[p];
        if (MODE == 2 && !(dist2_rn(px, py, pz, x, y, z) < r2d)) continue;
        cu.add(x, y, z);
    }
    kf_store_normal(cu, g.sidx[mypos], out);
}

