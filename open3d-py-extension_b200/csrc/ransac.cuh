// ransac.cuh — segment_plane on the device.
// Replaces pcd.segment_plane(distance_threshold, ransac_n, num_iterations, probability)
// (reference PointCloud.py:75-77; Open3D PointCloud::SegmentPlane).
//   sample  : Philox4x32-10 counter RNG on the device (replayable on the CPU bit for bit),
//             or a std::mt19937 replay of Open3D's RandomSampler, or caller-given triples
//   model   : one thread per hypothesis, FP64, unfused (ComputeTrianglePlane / GetPlaneFromPoints)
//   score   : |((a*x+b*y)+c*z)+d| < thr in FP64 for every (point, hypothesis) pair; a block
//             owns a slab of points and 32 hypotheses, every thread keeps its own running
//             (count, sum) column per hypothesis in shared memory, so there is no atomic and
//             the sums are reproducible from run to run
//   select  : the host replays Open3D's serial best-so-far / early-exit rule over the scores
//   final   : ordered compaction of the winner's inliers + two reduction passes for the refit
#pragma once
#include "knn_host.cuh"
#include <random>

#define PHILOX_TAG 0x52414E53u /* "RANS" */

__host__ __device__ inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1)
{
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

// counter = (draw, 0, iteration, TAG), key = seed; index = (c1:c0) mod n; a duplicate inside the
// sample is rejected by moving on to the next draw (Open3D RandomSampler's rejection rule)
__global__ void philox_sample_kernel(uint64_t seed, int64_t n, int rn, int H, int64_t *__restrict__ out)
{
    int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    int got = 0;
    uint32_t draw = 0;
    int64_t *row = out + (int64_t)h * rn;
    while (got < rn) {
        uint32_t c[4] = {draw++, 0u, (uint32_t)h, PHILOX_TAG};
        philox4x32_10(c, (uint32_t)seed, (uint32_t)(seed >> 32));
        uint64_t r = ((uint64_t)c[1] << 32) | c[0];
        int64_t cand = (int64_t)(r % (uint64_t)n);
        bool dup = false;
        for (int j = 0; j < got; ++j) dup |= row[j] == cand;
        if (!dup) row[got++] = cand;
    }
}

__device__ __forceinline__ void plane_from_sample(const double *__restrict__ x, const double *__restrict__ y,
                                                  const double *__restrict__ z, const int64_t *__restrict__ s,
                                                  int rn, double *pl)
{
    if (rn == 3) {
        double p0[3] = {x[s[0]], y[s[0]], z[s[0]]};
        double e0[3] = {x[s[1]] - p0[0], y[s[1]] - p0[1], z[s[1]] - p0[2]};
        double e1[3] = {x[s[2]] - p0[0], y[s[2]] - p0[1], z[s[2]] - p0[2]};
        double abc[3];
        cross3d(e0, e1, abc);
        double nn = sqrt(dot3d(abc, abc));
        if (nn == 0.0) { pl[0] = pl[1] = pl[2] = pl[3] = 0.0; return; }
        abc[0] /= nn; abc[1] /= nn; abc[2] /= nn;
        pl[0] = abc[0]; pl[1] = abc[1]; pl[2] = abc[2];
        pl[3] = -dot3d(abc, p0);
        return;
    }
    double c[3] = {0, 0, 0};
    for (int j = 0; j < rn; ++j) { c[0] += x[s[j]]; c[1] += y[s[j]]; c[2] += z[s[j]]; }
    double m = (double)rn;
    c[0] /= m; c[1] /= m; c[2] /= m;
    double xx = 0, xy = 0, xz = 0, yy = 0, yz = 0, zz = 0;
    for (int j = 0; j < rn; ++j) {
        double rx = x[s[j]] - c[0], ry = y[s[j]] - c[1], rz = z[s[j]] - c[2];
        xx += rx * rx; xy += rx * ry; xz += rx * rz;
        yy += ry * ry; yz += ry * rz; zz += rz * rz;
    }
    double det_x = yy * zz - yz * yz, det_y = xx * zz - xz * xz, det_z = xx * yy - xy * xy;
    double abc[3];
    if (det_x > det_y && det_x > det_z) { abc[0] = det_x; abc[1] = xz * yz - xy * zz; abc[2] = xy * yz - xz * yy; }
    else if (det_y > det_z)             { abc[0] = xz * yz - xy * zz; abc[1] = det_y; abc[2] = xy * xz - yz * xx; }
    else                                { abc[0] = xy * yz - xz * yy; abc[1] = xy * xz - yz * xx; abc[2] = det_z; }
    double nn = sqrt(dot3d(abc, abc));
    if (nn == 0.0) { pl[0] = pl[1] = pl[2] = pl[3] = 0.0; return; }
    abc[0] /= nn; abc[1] /= nn; abc[2] /= nn;
    pl[0] = abc[0]; pl[1] = abc[1]; pl[2] = abc[2];
    pl[3] = -dot3d(abc, c);
}

__global__ void plane_model_kernel(const double *__restrict__ x, const double *__restrict__ y,
                                   const double *__restrict__ z, const int64_t *__restrict__ samples, int rn,
                                   int H, double *__restrict__ planes)
{
    int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    double pl[4];
    plane_from_sample(x, y, z, samples + (int64_t)h * rn, rn, pl);
    planes[4 * h] = pl[0]; planes[4 * h + 1] = pl[1]; planes[4 * h + 2] = pl[2]; planes[4 * h + 3] = pl[3];
}

// ---- scoring -----------------------------------------------------------------------------------
// (Measured and not kept: an FP32 screen, t = a x + b y + c z + d in packed FP32 with a rigorous error margin, so that only
// points within thr + margin take the FP64 evaluation.  7 % of the point x hypothesis pairs of the 10 M scan are inliers, but
// 38 % of the 32-point warp steps touch one and a thread's four points share the branch: 1.13 ms against 0.97 ms.)
#define SC_THREADS 256
#ifndef SC_PPT
#define SC_PPT 4
#endif
#define SC_HB 32   // hypotheses per block (blockIdx.y selects the chunk)

__device__ __forceinline__ double plane_dist(double a, double b, double c, double d, double x, double y, double z)
{
    return fabs(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(a, x), __dmul_rn(b, y)), __dmul_rn(c, z)), d));
}

__global__ void __launch_bounds__(SC_THREADS)
score_kernel(const double *__restrict__ x, const double *__restrict__ y, const double *__restrict__ z, int64_t n,
             const double *__restrict__ planes, int h0, int H, int hper /* hypotheses per blockIdx.y, <= SC_HB */, double thr,
             unsigned long long *__restrict__ part_cnt /*[gridDim.x][Hbatch]*/, double *__restrict__ part_err)
{
    extern __shared__ unsigned char sc_smem[];
    double *serr = (double *)sc_smem;                                    // [SC_HB][SC_THREADS]
    uint32_t *scnt = (uint32_t *)(sc_smem + SC_HB * SC_THREADS * 8);     // [SC_HB][SC_THREADS]
    __shared__ double spl[SC_HB][4];
    const int tid = threadIdx.x;
    const int hbase = h0 + blockIdx.y * hper;
    const int nh = H - hbase < hper ? (H - hbase > 0 ? H - hbase : 0) : hper;
    if (tid < SC_HB * 4) {
        int hh = tid >> 2;
        spl[hh][tid & 3] = hh < nh ? planes[4 * (hbase + hh) + (tid & 3)] : 0.0;
    }
    for (int hh = 0; hh < SC_HB; ++hh) { serr[hh * SC_THREADS + tid] = 0.0; scnt[hh * SC_THREADS + tid] = 0u; }
    __syncthreads();
    // static slab of points per block: reproducible sums
    const int64_t tile = (int64_t)SC_THREADS * SC_PPT;
    const int64_t tiles = (n + tile - 1) / tile;
    const int64_t per = (tiles + gridDim.x - 1) / gridDim.x;
    const int64_t t0 = (int64_t)blockIdx.x * per, t1 = t0 + per < tiles ? t0 + per : tiles;
    for (int64_t t = t0; t < t1; ++t) {
        double px[SC_PPT], py[SC_PPT], pz[SC_PPT];
        bool ok[SC_PPT];
#pragma unroll
        for (int k = 0; k < SC_PPT; ++k) {
            int64_t i = t * tile + (int64_t)k * SC_THREADS + tid;
            ok[k] = i < n;
            px[k] = ok[k] ? x[i] : 0.0; py[k] = ok[k] ? y[i] : 0.0; pz[k] = ok[k] ? z[i] : 0.0;
        }
        for (int hh = 0; hh < nh; ++hh) {
            double a = spl[hh][0], b = spl[hh][1], c = spl[hh][2], d = spl[hh][3];
            double e = 0.0;
            uint32_t cn = 0;
#pragma unroll
            for (int k = 0; k < SC_PPT; ++k) {
                double dist = plane_dist(a, b, c, d, px[k], py[k], pz[k]);
                bool in = ok[k] && dist < thr;
                if (in) { e = __dadd_rn(e, dist); cn++; }
            }
            serr[hh * SC_THREADS + tid] = __dadd_rn(serr[hh * SC_THREADS + tid], e);
            scnt[hh * SC_THREADS + tid] += cn;
        }
    }
    __syncthreads();
    // column totals in a fixed order: warp w owns hypotheses w*4 .. w*4+3
    const int lane = lane_id(), w = tid >> 5;
    const int Hb = gridDim.y * SC_HB;
    for (int hh = w * (SC_HB / 8); hh < (w + 1) * (SC_HB / 8); ++hh) {
        double e = 0.0;
        unsigned long long cn = 0;
        for (int j = 0; j < SC_THREADS / 32; ++j) {
            e = __dadd_rn(e, serr[hh * SC_THREADS + j * 32 + lane]);
            cn += scnt[hh * SC_THREADS + j * 32 + lane];
        }
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            e = __dadd_rn(e, shfl_xor_d(e, s));
            cn += __shfl_xor_sync(FULL, cn, s);
        }
        if (lane == 0) {
            part_cnt[(int64_t)blockIdx.x * Hb + blockIdx.y * SC_HB + hh] = cn;
            part_err[(int64_t)blockIdx.x * Hb + blockIdx.y * SC_HB + hh] = e;
        }
    }
}

// sum the per-block partials in block order
__global__ void score_final_kernel(const unsigned long long *__restrict__ part_cnt, const double *__restrict__ part_err,
                                   int blocks, int Hb, int per, long long *__restrict__ cnt, double *__restrict__ err)
{
    int h = blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= Hb || h % SC_HB >= per) return;
    unsigned long long c = 0;
    double e = 0.0;
    for (int b = 0; b < blocks; ++b) { c += part_cnt[(int64_t)b * Hb + h]; e = __dadd_rn(e, part_err[(int64_t)b * Hb + h]); }
    const int o = (h / SC_HB) * per + h % SC_HB;        // columns of a chunk are packed `per` to a block row
    cnt[o] = (long long)c;
    err[o] = e;
}

struct PlanePred {
    const double *x, *y, *z;
    double a, b, c, d, thr;
    __device__ bool operator()(int64_t i) const { return plane_dist(a, b, c, d, x[i], y[i], z[i]) < thr; }
};

// mode 0: sums of coordinates; mode 1: the six centred second moments
#define RF_THREADS 256
__global__ void __launch_bounds__(RF_THREADS)
refit_sum_kernel(const int32_t *__restrict__ inl, int64_t m, const double *__restrict__ x, const double *__restrict__ y,
                 const double *__restrict__ z, double cx, double cy, double cz, int mode, double *__restrict__ part)
{
    double s[6] = {0, 0, 0, 0, 0, 0};
    int64_t per = (m + gridDim.x - 1) / gridDim.x;
    int64_t j0 = (int64_t)blockIdx.x * per, j1 = j0 + per < m ? j0 + per : m;
    for (int64_t j = j0 + threadIdx.x; j < j1; j += RF_THREADS) {
        int32_t i = inl[j];
        if (mode == 0) { s[0] += x[i]; s[1] += y[i]; s[2] += z[i]; }
        else {
            double rx = x[i] - cx, ry = y[i] - cy, rz = z[i] - cz;
            s[0] += rx * rx; s[1] += rx * ry; s[2] += rx * rz; s[3] += ry * ry; s[4] += ry * rz; s[5] += rz * rz;
        }
    }
    __shared__ double sh[RF_THREADS / 32][6];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) s[k] += shfl_xor_d(s[k], d);
    }
    if (lane_id() == 0) for (int k = 0; k < 6; ++k) sh[threadIdx.x >> 5][k] = s[k];
    __syncthreads();
    if (threadIdx.x < 6) {
        double v = 0.0;
        for (int w = 0; w < RF_THREADS / 32; ++w) v += sh[w][threadIdx.x];
        part[blockIdx.x * 6 + threadIdx.x] = v;
    }
}

static int segment_plane_impl(o3dpp_ctx *c, double thr, int rn, int H, double prob, int sampler, uint64_t seed,
                              const int64_t *given, double *plane4, int64_t *n_inliers, double *best_plane4,
                              int32_t *best_iter_out, int64_t *counts_out, double *errs_out)
{
    int64_t n = c->n;
    if (rn < 3) return set_err(O3DPP_E_RANSAC_N, "ransac_n should be set to higher than or equal to 3.");
    if (n < rn) return set_err(O3DPP_E_FEW_POINTS, "There must be at least 'ransac_n' points.");
    if (!(prob > 0.0) || prob > 1.0) return set_err(O3DPP_E_PROB, "Probability must be > 0 and <= 1.0");
    if (H < 0) return set_err(O3DPP_E_ARG, "num_iterations must be >= 0");
    if (sampler == O3DPP_SAMPLER_GIVEN && H > 0 && !given) return set_err(O3DPP_E_ARG, "samples required");
    if (sampler < 0 || sampler > 2) return set_err(O3DPP_E_ARG, "unknown sampler");
    c->r_ninl = -1;
    c->r_H = H; c->r_rn = rn;
    const double *X = c->pts.x(), *Y = c->pts.y(), *Z = c->pts.z();
    cudaEventRecord(c->ev[0], c->stream);
    // ---- hypothesis sequence
    size_t sbytes = (size_t)(H > 0 ? H : 1) * rn * sizeof(int64_t);
    O3_TRY(c->r_samples.ensure(sbytes));
    c->r_samples_host.assign((size_t)H * rn, 0);
    if (H > 0) {
        if (sampler == O3DPP_SAMPLER_PHILOX) {
            LAUNCH(c, philox_sample_kernel, (H + 127) / 128, 128, 0, seed, n, rn, H, c->r_samples.as<int64_t>());
            O3_TRY(check_launch("philox_sample"));
        } else {
            if (sampler == O3DPP_SAMPLER_MT19937) {
                std::mt19937 eng((uint32_t)seed);
                for (int h = 0; h < H; ++h) {
                    int got = 0;
                    int64_t *row = c->r_samples_host.data() + (size_t)h * rn;
                    while (got < rn) {
                        int64_t cand = (int64_t)((uint64_t)eng() % (uint64_t)n);
                        bool dup = false;
                        for (int j = 0; j < got; ++j) dup |= row[j] == cand;
                        if (!dup) row[got++] = cand;
                    }
                }
            } else {
                for (size_t i = 0; i < (size_t)H * rn; ++i) {
                    if (given[i] < 0 || given[i] >= n) return set_err(O3DPP_E_ARG, "sample index out of range");
                    c->r_samples_host[i] = given[i];
                }
            }
            CU_TRY(cudaMemcpyAsync(c->r_samples.p, c->r_samples_host.data(), sbytes, cudaMemcpyHostToDevice, c->stream));
        }
    }
    // ---- models
    O3_TRY(c->r_planes.ensure((size_t)(H + 64) * 4 * sizeof(double)));
    std::vector<double> planes((size_t)H * 4);
    if (H > 0) {
        LAUNCH(c, plane_model_kernel, (H + 127) / 128, 128, 0, X, Y, Z, c->r_samples.as<int64_t>(), rn, H, c->r_planes.as<double>());
        O3_TRY(check_launch("plane_model"));
        CU_TRY(cudaMemcpyAsync(planes.data(), c->r_planes.p, planes.size() * 8, cudaMemcpyDeviceToHost, c->stream));
        if (sampler == O3DPP_SAMPLER_PHILOX)
            CU_TRY(cudaMemcpyAsync(c->r_samples_host.data(), c->r_samples.p, sbytes, cudaMemcpyDeviceToHost, c->stream));
    }
    cudaEventRecord(c->ev[1], c->stream);
    CU_TRY(cudaStreamSynchronize(c->stream));
    // ---- score in batches, replaying Open3D's serial selection between batches
    const int BATCH = 64;
    int gx = c->sm_count;
    {
        int64_t tiles = (n + SC_THREADS * SC_PPT - 1) / (SC_THREADS * SC_PPT);
        if (tiles < gx) gx = (int)tiles;
        if (gx < 1) gx = 1;
    }
    O3_TRY(c->r_part.ensure((size_t)gx * BATCH * 16 + 256));
    O3_TRY(c->r_cnt.ensure(BATCH * 8 + 64));
    O3_TRY(c->r_err.ensure(BATCH * 8 + 64));
    unsigned long long *pc = c->r_part.as<unsigned long long>();
    double *pe = (double *)(pc + (size_t)gx * BATCH);
    size_t smem = (size_t)SC_HB * SC_THREADS * 12;
    CU_TRY(cudaFuncSetAttribute(score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    double best_fit = 0.0, best_rmse = 0.0, best_pl[4] = {0, 0, 0, 0};
    int best_iter = -1;
    double break_iteration = 1.7976931348623157e308;
    int64_t iteration_count = 0;
    int scored = 0;
    if (counts_out) for (int h = 0; h < H; ++h) { counts_out[h] = -1; if (errs_out) errs_out[h] = 0.0; }
    long long hc[BATCH];
    double he[BATCH];
    for (int h0 = 0, hb = 0; h0 < H; h0 += hb) {
        if ((double)iteration_count > break_iteration) break;
        hb = H - h0 < BATCH ? H - h0 : BATCH;
        // the serial loop stops after iteration break_iteration: do not score past it (the bound only ever drops)
        const double left = break_iteration + 1.0 - (double)iteration_count;
        if (left < (double)hb) hb = left < 1.0 ? 1 : (int)left;
        int gy = (hb + SC_HB - 1) / SC_HB;
        const int per = (hb + gy - 1) / gy;                 // even split: the blocks of a wave finish together
        dim3 grid(gx, gy);
        score_kernel<<<grid, SC_THREADS, smem, c->stream>>>(X, Y, Z, n, c->r_planes.as<double>(), h0, h0 + hb, per, thr, pc, pe);
        c->launches++;
        LAUNCH(c, score_final_kernel, 1, 64, 0, pc, pe, gx, gy * SC_HB, per, c->r_cnt.as<long long>(), c->r_err.as<double>());
        O3_TRY(check_launch("score"));
        CU_TRY(cudaMemcpyAsync(hc, c->r_cnt.p, hb * 8, cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(cudaMemcpyAsync(he, c->r_err.p, hb * 8, cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(cudaStreamSynchronize(c->stream));
        scored += hb;
        for (int b = 0; b < hb; ++b) {
            int h = h0 + b;
            if ((double)iteration_count > break_iteration) continue;
            const double *pl = &planes[(size_t)h * 4];
            if (pl[0] == 0.0 && pl[1] == 0.0 && pl[2] == 0.0 && pl[3] == 0.0) { if (counts_out) counts_out[h] = -2; continue; }
            if (counts_out) { counts_out[h] = hc[b]; if (errs_out) errs_out[h] = he[b]; }
            double fit, rmse;
            if (hc[b] == 0) { fit = 0.0; rmse = 0.0; }
            else { fit = (double)hc[b] / (double)n; rmse = he[b] / sqrt((double)hc[b]); }
            if (fit > best_fit || (fit == best_fit && rmse < best_rmse)) {
                best_fit = fit; best_rmse = rmse; best_iter = h;
                memcpy(best_pl, pl, sizeof best_pl);
                if (fit < 1.0) {
                    double bi = log(1.0 - prob) / log(1.0 - pow(fit, (double)rn));
                    if (!(bi >= 0.0)) break_iteration = 1.7976931348623157e308;
                    else break_iteration = floor(bi < (double)H ? bi : (double)H);
                } else break_iteration = 0.0;
            }
            iteration_count++;
        }
    }
    cudaEventRecord(c->ev[2], c->stream);
    // ---- final inliers of the winner (ascending index) and refit
    int64_t m = 0;
    bool zero = best_pl[0] == 0.0 && best_pl[1] == 0.0 && best_pl[2] == 0.0 && best_pl[3] == 0.0;
    O3_TRY(c->r_inl.ensure((size_t)(n + 64) * 4));
    if (!zero) {
        PlanePred pp{X, Y, Z, best_pl[0], best_pl[1], best_pl[2], best_pl[3], thr};
        O3_TRY(compact_indices(c, pp, n, c->r_inl.as<int32_t>(), &m));
    }
    double pl[4] = {0, 0, 0, 0};
    if (m > 0) {
        int g = grid_for(c, m, RF_THREADS * 8, 2);
        double *part = c->d_small.as<double>() + 16;
        double *hpart = (double *)c->h_small;
        double cen[3] = {0, 0, 0}, mom[6] = {0, 0, 0, 0, 0, 0};
        LAUNCH(c, refit_sum_kernel, g, RF_THREADS, 0, c->r_inl.as<int32_t>(), m, X, Y, Z, 0.0, 0.0, 0.0, 0, part);
        CU_TRY(cudaMemcpyAsync(hpart, part, (size_t)g * 6 * 8, cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(cudaStreamSynchronize(c->stream));
        for (int b = 0; b < g; ++b) for (int k = 0; k < 3; ++k) cen[k] += hpart[b * 6 + k];
        for (int k = 0; k < 3; ++k) cen[k] /= (double)m;
        LAUNCH(c, refit_sum_kernel, g, RF_THREADS, 0, c->r_inl.as<int32_t>(), m, X, Y, Z, cen[0], cen[1], cen[2], 1, part);
        O3_TRY(check_launch("refit"));
        CU_TRY(cudaMemcpyAsync(hpart, part, (size_t)g * 6 * 8, cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(cudaStreamSynchronize(c->stream));
        for (int b = 0; b < g; ++b) for (int k = 0; k < 6; ++k) mom[k] += hpart[b * 6 + k];
        double xx = mom[0], xy = mom[1], xz = mom[2], yy = mom[3], yz = mom[4], zz = mom[5];
        double det_x = yy * zz - yz * yz, det_y = xx * zz - xz * xz, det_z = xx * yy - xy * xy;
        double abc[3];
        if (det_x > det_y && det_x > det_z) { abc[0] = det_x; abc[1] = xz * yz - xy * zz; abc[2] = xy * yz - xz * yy; }
        else if (det_y > det_z)             { abc[0] = xz * yz - xy * zz; abc[1] = det_y; abc[2] = xy * xz - yz * xx; }
        else                                { abc[0] = xy * yz - xz * yy; abc[1] = xy * xz - yz * xx; abc[2] = det_z; }
        double nn = sqrt((abc[0] * abc[0] + abc[1] * abc[1]) + abc[2] * abc[2]);
        if (nn != 0.0) {
            abc[0] /= nn; abc[1] /= nn; abc[2] /= nn;
            pl[0] = abc[0]; pl[1] = abc[1]; pl[2] = abc[2];
            pl[3] = -((abc[0] * cen[0] + abc[1] * cen[1]) + abc[2] * cen[2]);
        }
    }
    cudaEventRecord(c->ev[3], c->stream);
    CU_TRY(cudaStreamSynchronize(c->stream));
    c->r_ninl = m;
    if (plane4) memcpy(plane4, pl, sizeof pl);
    if (n_inliers) *n_inliers = m;
    if (best_plane4) memcpy(best_plane4, best_pl, sizeof best_pl);
    if (best_iter_out) *best_iter_out = best_iter;
    float t0 = 0, t1 = 0, t2 = 0;
    cudaEventElapsedTime(&t0, c->ev[0], c->ev[1]);
    cudaEventElapsedTime(&t1, c->ev[1], c->ev[2]);
    cudaEventElapsedTime(&t2, c->ev[2], c->ev[3]);
    c->timings[0] = t0; c->timings[1] = t1; c->timings[2] = t2; c->timings[7] = scored;
    return 0;
}
