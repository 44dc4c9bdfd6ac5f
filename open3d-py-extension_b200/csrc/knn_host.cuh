// knn_host.cuh — host side of the neighbour search: picks the kernel tier chain and launches it.
#pragma once
#include "knn_lane.cuh"

// ---- host side -----------------------------------------------------------------------------------
struct CapPending {
    const float *cap;
    __device__ __forceinline__ bool operator()(int64_t i) const { return cap[i] >= 0.0f; }
};

// tier chain of the lane-per-query path (knn_lane.cuh): tier 1 over every cell group, then the queries it
// could not certify, compacted in curve order, through the FP64 range query and the exact list search
template <int MODE>
static int launch_lane(o3dpp_ctx *c, const GridDesc &gd, int K, double radius, KnnOut out)
{
    int64_t n = gd.n;
    O3_TRY(c->g_ncnt.ensure((size_t)n * 4 + 64));
    O3_TRY(c->g_nbr.ensure((size_t)n * 4 + 64));
    float *capbuf = c->g_ncnt.as<float>();
    int32_t *list = c->g_nbr.as<int32_t>();
    // neighbour positions [K][pitch]: tier 1 -> knn_finish_kernel; in keep mode every tier records its sets
    const int64_t pitch = plane_stride(n);
    O3_TRY(c->g_nbrpos.ensure((size_t)pitch * K * 4 + 64));
    out.nbr = c->g_nbrpos.as<uint32_t>();
    out.nbr_pitch = pitch;
    out.keep = c->tune.keep ? 1 : 0;
    out.tier = nullptr;
    c->nbr_valid = false;
    if (out.keep) {
        O3_TRY(c->g_tier.ensure((size_t)n + 64));
        out.tier = c->g_tier.as<unsigned char>();
        CU_TRY(cudaMemsetAsync(out.tier, 0, (size_t)n, c->stream));
    }
    // rho = staging radius / home-box diagonal (0.42: measured optimum on the LiDAR-like scans, all tiers, 10 M: 0.44 -> 11.43 ms,
    // 0.42 -> 11.30, 0.40 -> 11.4).  A list longer than gsplit,
    // or candidates that overflow the staging area, retry with the radius x shrink, up to max_shrink times
    // (gsplit at 10M, all three tiers: 48 -> 13.40 ms, 64 -> 13.23, 80 -> 13.30, 96 -> 13.42, 128 -> 13.47)
    const KnnTune &t = c->tune;
    const float rho = t.rho > 0.f ? t.rho : 0.42f;
    const int max_shrink = t.max_shrink >= 0 && t.max_shrink <= 12 ? t.max_shrink : 12;
    const int gsplit = t.gsplit >= 16 && t.gsplit <= KS_GMAX - 64 ? t.gsplit : 64;
    const float shrink = t.shrink > 0.1f && t.shrink < 0.99f ? t.shrink : 0.7f;
    // group lists (knn_groups_kernel) -> selection (knn_select_kernel)
    KsGroups gr;
    gr.pitch = ((gsplit + 31) / 32) * 32;
    const int64_t ngr = gd.cnt[0];
    O3_TRY(c->g_glist.ensure((size_t)ngr * gr.pitch * 4 + 64));
    O3_TRY(c->g_gcnt.ensure((size_t)ngr * 3 + 64));
    gr.list = c->g_glist.as<uint32_t>();
    gr.cnt = c->g_gcnt.as<unsigned short>();
    gr.tries = c->g_gcnt.as<unsigned char>() + (size_t)ngr * 2;
    const size_t smem = sizeof(KsMem);
    auto k1 = knn_select_kernel<MODE, KS_W>;
    CU_TRY(cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEventRecord(c->ev[3], c->stream);
    LAUNCH(c, knn_groups_kernel<MODE>, (int)((ngr + KA_WARPS - 1) / KA_WARPS), KA_WARPS * 32, 0, gd, radius, rho, max_shrink, gsplit, shrink, gr);
    O3_TRY(check_launch("knn_groups"));
    cudaEventRecord(c->ev[7], c->stream);
    LAUNCH(c, k1, (int)ngr, KS_W * 32, smem, gd, K, radius, rho, max_shrink, shrink, gr, out, capbuf);
    O3_TRY(check_launch("knn_select"));
    cudaEventRecord(c->ev[4], c->stream);
    if (out.nx) {
        LAUNCH(c, knn_finish_kernel<MODE>, (int)((n + 127) / 128), 128, 0, gd, K, radius, out, capbuf);
        O3_TRY(check_launch("knn_finish"));
    }
    cudaEventRecord(c->ev[6], c->stream);
    int64_t m2 = 0, m3 = 0;
    CapPending pred{capbuf};
    O3_TRY(compact_indices(c, pred, n, list, &m2));
    if (m2 > 0) {
        auto k2 = knn_range_kernel<MODE>;
        // queries per warp: 32 when the failed queries fill the machine many times over, fewer (down to 2) when they do not
        int qpw = (int)(m2 / ((int64_t)c->sm_count * 32));
        qpw = qpw < 2 ? 2 : (qpw > 32 ? 32 : qpw);
        int64_t warps2 = (m2 + qpw - 1) / qpw;
        O3_TRY(c->g_park.ensure((size_t)m2 * 11 * sizeof(double) + 256));
        LAUNCH(c, k2, (int)((warps2 + K2_WARPS - 1) / K2_WARPS), K2_WARPS * 32, 0, gd, K, radius, out, list, (int)m2, capbuf, qpw, c->g_park.as<double>());
        LAUNCH(c, knn_range_finish_kernel<MODE>, (int)((m2 + 127) / 128), 128, 0, gd, out, list, (int)m2, capbuf, c->g_park.as<double>());
        O3_TRY(check_launch("knn_range"));
        cudaEventRecord(c->ev[5], c->stream);
        O3_TRY(compact_indices(c, pred, n, list, &m3));
        if (m3 > 0) {
            auto k3 = knn_exact_kernel<MODE>;
            LAUNCH(c, k3, (int)((m3 + 3) / 4), 128, 0, gd, K, radius, out, list, (int)m3, capbuf);
            O3_TRY(check_launch("knn_exact"));
        }
    } else cudaEventRecord(c->ev[5], c->stream);
    c->timings[4] = (double)m2;
    c->timings[5] = (double)m3;
    c->kth_valid = true;
    if (out.keep) { c->nbr_valid = true; c->nbr_K = K; c->nbr_mode = MODE; c->nbr_pitch = pitch; c->nbr_radius = radius; }
    return 0;
}

template <int MODE, int S, int WARPS>
static int launch_search(o3dpp_ctx *c, const GridDesc &gd, int K, double radius, const KnnOut &out)
{
    size_t smem = (size_t)WARPS * 32 * (32 * S + 1) * sizeof(int);
    auto kern = knn_normals_kernel<MODE, S, WARPS>;
    if (smem > 48 * 1024) CU_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t blocks = (gd.cnt[0] + WARPS - 1) / WARPS;
    LAUNCH(c, kern, (int)blocks, WARPS * 32, smem, gd, K, radius, out);
    return check_launch("knn_normals");
}

template <int MODE>
static int launch_search_mode(o3dpp_ctx *c, const GridDesc &gd, int K, double radius, const KnnOut &out)
{
    if (MODE == 1 || K <= 32) return launch_search<MODE, 1, 4>(c, gd, K, radius, out);
    if (K <= 64) return launch_search<MODE, 2, 4>(c, gd, K, radius, out);
    return launch_search<MODE, 4, 2>(c, gd, K, radius, out);
}

static int search_impl(o3dpp_ctx *c, int mode, int K, double radius, bool want_normals, int32_t *h_idx,
                       double *h_d2, int32_t *h_cnt)
{
    int64_t n = c->n;
    if (mode < 0 || mode > 2) return set_err(O3DPP_E_ARG, "unknown search mode");
    if (mode != O3DPP_SEARCH_RADIUS) {
        if (K <= 0) return set_err(O3DPP_E_ARG, "knn must be positive");
        if (K > O3DPP_MAX_KNN) return set_err(O3DPP_E_UNSUPPORTED, "knn larger than O3DPP_MAX_KNN");
    } else {
        if (h_idx || h_d2) return set_err(O3DPP_E_UNSUPPORTED, "radius search returns counts only");
        K = 1;
    }
    if (mode != O3DPP_SEARCH_KNN && !(radius >= 0.0)) return set_err(O3DPP_E_ARG, "radius must be >= 0");
    if (n == 0) return 0;
    KnnOut out;
    memset(&out, 0, sizeof out);
    if (want_normals) {
        out.has_old_normals = c->nrm.present ? 1 : 0;
        if (!c->nrm.present) O3_TRY(attr_alloc(&c->nrm, n));
        out.nx = c->nrm.x(); out.ny = c->nrm.y(); out.nz = c->nrm.z();
    }
    if (h_idx) { O3_TRY(c->g_nbr.ensure((size_t)n * K * 4 + 64)); out.idx = c->g_nbr.as<int32_t>(); }
    if (h_d2) { O3_TRY(c->g_nd2.ensure((size_t)n * K * 8 + 64)); out.d2 = c->g_nd2.as<double>(); }
    if (h_cnt) { O3_TRY(c->g_ncnt.ensure((size_t)n * 4 + 64)); out.cnt = c->g_ncnt.as<int32_t>(); }
    if ((h_d2 || h_cnt) && !h_idx) return set_err(O3DPP_E_ARG, "d2 / cnt need idx");

    if (c->want_stats) {
        out.stats = (unsigned long long *)(c->d_small.as<unsigned char>() + 56 * 1024);
        CU_TRY(cudaMemsetAsync(out.stats, 0, 512, c->stream));
    }
    GridDesc gd;
    c->kth_valid = false;
    cudaEventRecord(c->ev[0], c->stream);
    O3_TRY(grid_build(c, &gd));
    cudaEventRecord(c->ev[1], c->stream);
    int rc;
    // estimate_normals with k <= 32 takes the certified FP32 selection kernel; searches that must
    // return sorted (d2, index) lists, k > 32 and pure radius search use the FP64 list kernel
    const bool fast = want_normals && !h_idx && K <= 32 && mode != O3DPP_SEARCH_RADIUS && !c->tune.force_list;
    if (fast) rc = mode == 0 ? launch_lane<0>(c, gd, K, radius, out) : launch_lane<2>(c, gd, K, radius, out);
    else if (mode == 0) rc = launch_search_mode<0>(c, gd, K, radius, out);
    else if (mode == 1) rc = launch_search_mode<1>(c, gd, K, radius, out);
    else rc = launch_search_mode<2>(c, gd, K, radius, out);
    if (rc) return rc;
    cudaEventRecord(c->ev[2], c->stream);
    if (want_normals) c->nrm.present = true;
    O3_TRY(d2h(c, h_idx, c->g_nbr.p, (size_t)n * K * 4));
    O3_TRY(d2h(c, h_d2, c->g_nd2.p, (size_t)n * K * 8));
    O3_TRY(d2h(c, h_cnt, c->g_ncnt.p, (size_t)n * 4));
    if (c->want_stats) O3_TRY(d2h(c, c->stats, out.stats, 512));
    CU_TRY(cudaStreamSynchronize(c->stream));
    float a = 0, b = 0;
    cudaEventElapsedTime(&a, c->ev[0], c->ev[1]);
    cudaEventElapsedTime(&b, c->ev[1], c->ev[2]);
    c->timings[0] = a;
    c->timings[1] = b;
    c->timings[2] = c->timings[3] = c->timings[6] = c->timings[7] = 0.0;
    if (fast) {
        float t1 = 0, t2 = 0, t3 = 0, tf = 0;
        float tg = 0;
        cudaEventElapsedTime(&tg, c->ev[3], c->ev[7]);      // group lists, part of t1
        cudaEventElapsedTime(&t1, c->ev[3], c->ev[4]);
        cudaEventElapsedTime(&tf, c->ev[4], c->ev[6]);
        cudaEventElapsedTime(&t2, c->ev[6], c->ev[5]);
        cudaEventElapsedTime(&t3, c->ev[5], c->ev[2]);
        c->timings[2] = t1;     // tier 1 selection kernel
        c->timings[7] = tf;     // finish kernel (cumulants + eigenvector of the tier-1 answers)
        c->timings[3] = t2;     // compaction + tier 2
        c->timings[6] = t3;     // compaction + tier 3
        if (c->want_stats && c->stats[16]) {
            double nq = (double)n, nw = (double)c->stats[16];
            fprintf(stderr, "[o3dpp lane] n=%lld mode=%d K=%d tier1 %.3f ms (lists %.3f), finish %.3f ms, tier2 %.3f ms, tier3 %.3f ms (tier2 %lld, tier3 %lld queries) | per warp: parts %.3f candidates %.1f "
                    "groups-scanned %.1f restages %.3f | fail per query: gap %.4f rho %.4f boundary %.4f count %.4f warp %.4f | split warps %.3f, kcycles per warp: single %.1f split %.1f\n", (long long)n, mode, K,
                    t1, tg, tf, t2, t3, (long long)c->timings[4], (long long)c->timings[5], c->stats[25] / nw, c->stats[17] / nw, c->stats[18] / nw, c->stats[24] / nw,
                    c->stats[19] / nq, c->stats[20] / nq, c->stats[21] / nq, c->stats[22] / nq, c->stats[23] / nq,
                    c->stats[28] / nw, c->stats[26] / (nw - c->stats[28] + 1e-9) / 1e3, c->stats[27] / (c->stats[28] + 1e-9) / 1e3);
            fprintf(stderr, "[o3dpp lane] by staged candidates (buckets of 96): groups");
            for (int b = 0; b < 8; ++b) fprintf(stderr, " %llu", c->stats[40 + b]);
            fprintf(stderr, " | lanes failing the radius test");
            for (int b = 0; b < 8; ++b) fprintf(stderr, " %llu", c->stats[32 + b]);
            fprintf(stderr, "\n");
        }
    }
    if (c->want_stats) {
        double nq = (double)n, nw = (double)((n + 31) / 32);
        fprintf(stderr, "[o3dpp stats] n=%lld mode=%d K=%d ms=%.3f | per query: slow %.4f roundsA %.2f roundsC %.2f failwarp %.4f failA %.4f failC %.4f "
                "failgap %.4f failrho %.4f active-rows %.2f tier2 %.4f | per warp: rows %.2f groups-scanned %.2f\n", (long long)n, mode, K, b,
                c->stats[4] / nq, c->stats[6] / nq, c->stats[7] / nq, c->stats[8] / nq, c->stats[9] / nq, c->stats[10] / nq,
                c->stats[11] / nq, c->stats[12] / nq, c->stats[14] / nq, c->stats[15] / nq, c->stats[5] / nw, c->stats[13] / nw);
    }
    return 0;
}


// per point (original index): an upper bound of the squared distance to its k-th neighbour, as left by
// the lane path of the last estimate_normals (exact for queries answered by tiers 2 and 3).  The slab-sharded
// KNN uses it to tell which queries near a cut need a wider halo (open3dpypro/sharding.py).
__global__ void kth_scatter_kernel(const float *__restrict__ capbuf, const uint32_t *__restrict__ sidx, int64_t n, float *__restrict__ out)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        float v = capbuf[i];
        out[sidx[i]] = v < 0.f ? -v : __int_as_float(0x7f800000);
    }
}

static int fetch_kth_impl(o3dpp_ctx *c, float *host, float *dev = nullptr)
{
    if (!c->kth_valid) return set_err(O3DPP_E_STATE, "no k-th distance bounds: run o3dpp_estimate_normals (knn <= 32, KNN or hybrid) first");
    int64_t n = c->n;
    O3_TRY(c->g_nd2.ensure((size_t)n * 4 + 64));
    LAUNCH(c, kth_scatter_kernel, grid_for(c, n, 256, 16), 256, 0, c->g_ncnt.as<float>(), c->g_sidx.as<uint32_t>(), n, dev ? dev : c->g_nd2.as<float>());
    O3_TRY(check_launch("kth_scatter"));
    if (!dev) O3_TRY(d2h(c, host, c->g_nd2.p, (size_t)n * 4));
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}


// o3dpp_fetch_neighbors: the neighbour sets the tier chain of the last estimate_normals selected (keep mode), as ascending
// original indices.  queries = m original indices (host), nullptr = all points in order.
static int fetch_neighbors_impl(o3dpp_ctx *c, const int64_t *queries, int64_t m, int32_t *idx, int32_t *cnt, int8_t *tier)
{
    if (!c->nbr_valid) return set_err(O3DPP_E_STATE, "no neighbour sets: set O3DPP_OPT_KEEP_NEIGHBORS and run o3dpp_estimate_normals (knn <= 32, KNN or hybrid) first");
    if (!idx) return set_err(O3DPP_E_ARG, "idx is required");
    const int64_t n = c->n;
    if (!queries) m = n;
    if (m < 0) return set_err(O3DPP_E_ARG, "negative query count");
    if (m == 0) return 0;
    const int K = c->nbr_K;
    GridDesc gd;
    memset(&gd, 0, sizeof gd);
    gd.n = n;
    gd.sp4 = c->g_sp4.as<double4>();
    gd.sidx = c->g_sidx.as<uint32_t>();
    O3_TRY(c->g_inv.ensure((size_t)n * 4 + 64));
    LAUNCH(c, inverse_perm_kernel, grid_for(c, n, 256, 16), 256, 0, gd.sidx, n, c->g_inv.as<uint32_t>());
    O3_TRY(check_launch("inverse_perm"));
    // scratch: queries | idx | cnt | tier
    const size_t oq = 0, oi = oq + (size_t)m * 8, oc = oi + (size_t)m * K * 4, ot = oc + (size_t)m * 4, total = ot + (size_t)m + 64;
    O3_TRY(c->g_nd2.ensure(total));
    unsigned char *base = c->g_nd2.as<unsigned char>();
    if (queries) CU_TRY(cudaMemcpyAsync(base + oq, queries, (size_t)m * 8, cudaMemcpyHostToDevice, c->stream));
    LAUNCH(c, nbr_fetch_kernel, (int)((m + 127) / 128), 128, 0, gd, c->g_nbrpos.as<uint32_t>(), c->nbr_pitch, c->g_tier.as<unsigned char>(),
           c->g_inv.as<uint32_t>(), queries ? (const int64_t *)(base + oq) : nullptr, m, K, c->nbr_mode, c->nbr_radius,
           (int32_t *)(base + oi), (int32_t *)(base + oc), (int8_t *)(base + ot));
    O3_TRY(check_launch("nbr_fetch"));
    O3_TRY(d2h(c, idx, base + oi, (size_t)m * K * 4));
    O3_TRY(d2h(c, cnt, base + oc, (size_t)m * 4));
    O3_TRY(d2h(c, tier, base + ot, (size_t)m));
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}
