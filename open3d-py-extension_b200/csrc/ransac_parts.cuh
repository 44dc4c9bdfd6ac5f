// ransac_parts.cuh — the steps of segment_plane as separate entry points, for a scan that is sharded
// across GPUs by spatial slab: every rank scores ALL hypotheses on its own
// points, the per-hypothesis (count, error) pairs are summed across ranks by the host side
// (open3dpypro/sharding.py: all-reduce), the host replays Open3D's serial selection rule, and the
// final inlier pass + refit again run per rank with an all-reduce of the centroid sums and moments.
// Same kernels and the same FP64 expressions as the single-GPU path in ransac.cuh.
#pragma once
#include "ransac.cuh"

// H*rn indices in [0, n): the Philox sequence segment_plane(sampler = PHILOX, seed) would use on a cloud of n points
static int sample_philox_impl(o3dpp_ctx *c, uint64_t seed, int64_t n, int rn, int H, int64_t *out)
{
    if (rn < 1 || H < 0 || n < rn) return set_err(O3DPP_E_ARG, "bad sampling request");
    if (H == 0) return 0;
    size_t bytes = (size_t)H * rn * sizeof(int64_t);
    O3_TRY(c->r_samples.ensure(bytes));
    LAUNCH(c, philox_sample_kernel, (H + 127) / 128, 128, 0, seed, n, rn, H, c->r_samples.as<int64_t>());
    O3_TRY(check_launch("philox_sample"));
    O3_TRY(d2h(c, out, c->r_samples.p, bytes));
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

__global__ void iota_kernel(int64_t *out, int64_t m)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < m) out[i] = i;
}

// plane models of H samples given by their coordinates: pts = H*rn rows (x, y, z), host
static int planes_from_points_impl(o3dpp_ctx *c, const double *pts, int rn, int H, double *planes)
{
    if (rn < 3 || H < 0) return set_err(O3DPP_E_ARG, "bad model request");
    if (H == 0) return 0;
    int64_t m = (int64_t)H * rn;
    int64_t stride = plane_stride(m);
    std::vector<double> soa((size_t)stride * 3, 0.0);
    for (int64_t i = 0; i < m; ++i) { soa[i] = pts[3 * i]; soa[stride + i] = pts[3 * i + 1]; soa[2 * stride + i] = pts[3 * i + 2]; }
    O3_TRY(c->stage.ensure(soa.size() * 8));
    O3_TRY(c->r_samples.ensure((size_t)m * 8));
    O3_TRY(c->r_planes.ensure((size_t)(H + 64) * 4 * sizeof(double)));
    CU_TRY(cudaMemcpyAsync(c->stage.p, soa.data(), soa.size() * 8, cudaMemcpyHostToDevice, c->stream));
    LAUNCH(c, iota_kernel, (int)((m + 255) / 256), 256, 0, c->r_samples.as<int64_t>(), m);
    const double *X = c->stage.as<double>();
    LAUNCH(c, plane_model_kernel, (H + 127) / 128, 128, 0, X, X + stride, X + 2 * stride, c->r_samples.as<int64_t>(), rn, H,
           c->r_planes.as<double>());
    O3_TRY(check_launch("plane_model"));
    O3_TRY(d2h(c, planes, c->r_planes.p, (size_t)H * 4 * 8));
    CU_TRY(cudaStreamSynchronize(c->stream));
    return 0;
}

// (count, error sum) of every given plane over the first n_use points of the cloud (0 = all)
static int score_planes_impl(o3dpp_ctx *c, const double *planes, int H, double thr, int64_t n_use, int64_t *counts, double *errs)
{
    int64_t n = n_use > 0 && n_use < c->n ? n_use : c->n;
    if (H < 0) return set_err(O3DPP_E_ARG, "H must be >= 0");
    if (H == 0) return 0;
    const double *X = c->pts.x(), *Y = c->pts.y(), *Z = c->pts.z();
    O3_TRY(c->r_planes.ensure((size_t)(H + 64) * 4 * sizeof(double)));
    CU_TRY(cudaMemcpyAsync(c->r_planes.p, planes, (size_t)H * 32, cudaMemcpyHostToDevice, c->stream));
    const int BATCH = 64;
    int gx = c->sm_count;
    {
        int64_t tiles = (n + SC_THREADS * SC_PPT - 1) / (SC_THREADS * SC_PPT);
        if (tiles < gx) gx = (int)tiles;
        if (gx < 1) gx = 1;
    }
    O3_TRY(c->r_part.ensure((size_t)gx * BATCH * 16 + 256));
    int Hpad = ((H + BATCH - 1) / BATCH) * BATCH;
    O3_TRY(c->r_cnt.ensure((size_t)Hpad * 8 + 64));
    O3_TRY(c->r_err.ensure((size_t)Hpad * 8 + 64));
    unsigned long long *pc = c->r_part.as<unsigned long long>();
    double *pe = (double *)(pc + (size_t)gx * BATCH);
    size_t smem = (size_t)SC_HB * SC_THREADS * 12;
    CU_TRY(cudaFuncSetAttribute(score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEventRecord(c->ev[0], c->stream);
    for (int h0 = 0; h0 < H; h0 += BATCH) {
        int hb = H - h0 < BATCH ? H - h0 : BATCH;
        int gy = (hb + SC_HB - 1) / SC_HB;
        dim3 grid(gx, gy);
        score_kernel<<<grid, SC_THREADS, smem, c->stream>>>(X, Y, Z, n, c->r_planes.as<double>(), h0, H, SC_HB, thr, pc, pe);
        c->launches++;
        LAUNCH(c, score_final_kernel, 1, 64, 0, pc, pe, gx, gy * SC_HB, SC_HB, c->r_cnt.as<long long>() + h0, c->r_err.as<double>() + h0);
    }
    O3_TRY(check_launch("score_planes"));
    cudaEventRecord(c->ev[1], c->stream);
    O3_TRY(d2h(c, counts, c->r_cnt.p, (size_t)H * 8));
    O3_TRY(d2h(c, errs, c->r_err.p, (size_t)H * 8));
    CU_TRY(cudaStreamSynchronize(c->stream));
    float t = 0;
    cudaEventElapsedTime(&t, c->ev[0], c->ev[1]);
    c->timings[1] = t;
    return 0;
}

// inliers of one plane among the first n_use points: the list stays on the device (o3dpp_fetch_inliers /
// o3dpp_select_inliers), sum3 = coordinate sums of the inliers (for the global centroid)
static int plane_inliers_impl(o3dpp_ctx *c, const double *pl, double thr, int64_t n_use, int64_t *n_inliers, double *sum3)
{
    int64_t n = n_use > 0 && n_use < c->n ? n_use : c->n;
    const double *X = c->pts.x(), *Y = c->pts.y(), *Z = c->pts.z();
    c->r_ninl = -1;
    int64_t m = 0;
    O3_TRY(c->r_inl.ensure((size_t)(c->n + 64) * 4));
    bool zero = pl[0] == 0.0 && pl[1] == 0.0 && pl[2] == 0.0 && pl[3] == 0.0;
    if (!zero && n > 0) {
        PlanePred pp{X, Y, Z, pl[0], pl[1], pl[2], pl[3], thr};
        O3_TRY(compact_indices(c, pp, n, c->r_inl.as<int32_t>(), &m));
    }
    double s[3] = {0, 0, 0};
    if (m > 0) {
        int g = grid_for(c, m, RF_THREADS * 8, 2);
        double *part = c->d_small.as<double>() + 16;
        double *hpart = (double *)c->h_small;
        LAUNCH(c, refit_sum_kernel, g, RF_THREADS, 0, c->r_inl.as<int32_t>(), m, X, Y, Z, 0.0, 0.0, 0.0, 0, part);
        O3_TRY(check_launch("refit"));
        CU_TRY(cudaMemcpyAsync(hpart, part, (size_t)g * 6 * 8, cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(cudaStreamSynchronize(c->stream));
        for (int b = 0; b < g; ++b) for (int k = 0; k < 3; ++k) s[k] += hpart[b * 6 + k];
    }
    c->r_ninl = m;
    if (n_inliers) *n_inliers = m;
    if (sum3) memcpy(sum3, s, sizeof s);
    return 0;
}

// the six second moments of the current inlier list about `centroid` (xx, xy, xz, yy, yz, zz)
static int inlier_moments_impl(o3dpp_ctx *c, const double *cen, double *mom6)
{
    if (c->r_ninl < 0) return set_err(O3DPP_E_STATE, "no inlier list on the current cloud");
    double mom[6] = {0, 0, 0, 0, 0, 0};
    int64_t m = c->r_ninl;
    if (m > 0) {
        int g = grid_for(c, m, RF_THREADS * 8, 2);
        double *part = c->d_small.as<double>() + 16;
        double *hpart = (double *)c->h_small;
        LAUNCH(c, refit_sum_kernel, g, RF_THREADS, 0, c->r_inl.as<int32_t>(), m, c->pts.x(), c->pts.y(), c->pts.z(), cen[0], cen[1],
               cen[2], 1, part);
        O3_TRY(check_launch("refit"));
        CU_TRY(cudaMemcpyAsync(hpart, part, (size_t)g * 6 * 8, cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(cudaStreamSynchronize(c->stream));
        for (int b = 0; b < g; ++b) for (int k = 0; k < 6; ++k) mom[k] += hpart[b * 6 + k];
    }
    memcpy(mom6, mom, sizeof mom);
    return 0;
}
