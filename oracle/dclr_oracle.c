/*
 * dclr_oracle.c -- CPU restatement of the reference's point-set ops.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the *checker* for the CUDA path in deepclr_b200/csrc.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it; the product never does.
 *
 * Every function follows one reference kernel line by line (paths relative to /root/reference):
 *   oracle_fps         <- extern/Pointnet2.PyTorch/pointnet2/src/sampling_gpu.cu:94-209
 *                         (+ caller-side init pointnet2_utils.py:25-26: temp = 1e10, idx int32)
 *   oracle_fps_lowidx  <- same arithmetic, globally-lowest-index tie-break (the rule BASELINE.json states)
 *   oracle_gather      <- sampling_gpu.cu:8-24
 *   oracle_ball_query  <- ball_query_gpu.cu:9-45 (+ caller zero-fill pointnet2_utils.py:218)
 *   oracle_group       <- group_points_gpu.cu:47-66
 *   oracle_knn         <- torch-cluster 1.5.9 knn (requirements.txt:24; NOT vendored in the reference):
 *                         brute force, squared L2 accumulated x,y,z with FMA contraction, strict '>' insertion
 *                         (ties keep the lower candidate index), output [row;col] int64 global row ids, as
 *                         consumed at deepclr/models/deepclr.py:164-171.  PARITY UNPINNED for this op: the
 *                         third-party source is absent, the restatement follows its published algorithm.
 *
 * Floating point: nvcc (-fmad=true) contracts the reference's squared distance into
 *     t = dy*dy ; t = fma(dx,dx,t) ; d = fma(dz,dz,t)
 * (observed in the sm_100a SASS of both kernels).  We spell that order with fmaf and build
 * with -ffp-contract=off so gcc adds no contraction of its own.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

static inline float sqdist_pn2(float dx, float dy, float dz) {
    float t = dy * dy;
    t = fmaf(dx, dx, t);
    return fmaf(dz, dz, t);
}

/* largest power of two <= n, capped at 1024, at least 1: cuda_utils.h:10-14 */
static int opt_n_threads(int work_size) {
    int p = 1;
    while (p * 2 <= work_size && p * 2 <= 1024) p *= 2;
    return p;
}

/*
 * Reference-faithful FPS including the block-tree tie-break (sampling_gpu.cu:86-91,143-203):
 * thread t owns points t, t+bs, ...; keeps its first strict maximum; the tree keeps the lower slot on ties.
 * xyz: [B,N,3]; idx out: [B,M] int32.  temp is allocated here and initialised to 1e10 like the caller does.
 */
void oracle_fps(const float *xyz, int B, int N, int M, int32_t *idx_out) {
    if (M <= 0) return;
    int bs = opt_n_threads(N);
#pragma omp parallel for schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        const float *p = xyz + (size_t)b * N * 3;
        int32_t *out = idx_out + (size_t)b * M;
        float *temp = (float *)malloc(sizeof(float) * (size_t)(N > 0 ? N : 1));
        float *dists = (float *)malloc(sizeof(float) * (size_t)bs);
        int *dists_i = (int *)malloc(sizeof(int) * (size_t)bs);
        for (int k = 0; k < N; ++k) temp[k] = 1e10f;
        int old = 0;
        out[0] = old;
        for (int j = 1; j < M; ++j) {
            float x1 = p[old * 3 + 0], y1 = p[old * 3 + 1], z1 = p[old * 3 + 2];
            for (int t = 0; t < bs; ++t) {
                int besti = 0;
                float best = -1.0f;
                for (int k = t; k < N; k += bs) {
                    float d = sqdist_pn2(p[k * 3 + 0] - x1, p[k * 3 + 1] - y1, p[k * 3 + 2] - z1);
                    float d2 = d < temp[k] ? d : temp[k];
                    temp[k] = d2;
                    besti = d2 > best ? k : besti;
                    best = d2 > best ? d2 : best;
                }
                dists[t] = best;
                dists_i[t] = besti;
            }
            for (int s = bs / 2; s >= 1; s /= 2) {
                for (int t = 0; t < s; ++t) {
                    float v1 = dists[t], v2 = dists[t + s];
                    int i1 = dists_i[t], i2 = dists_i[t + s];
                    dists[t] = v1 > v2 ? v1 : v2;
                    dists_i[t] = v2 > v1 ? i2 : i1;
                }
            }
            old = dists_i[0];
            out[j] = old;
        }
        free(temp);
        free(dists);
        free(dists_i);
    }
}

/*
 * Same arithmetic, tie-break = globally lowest point index (BASELINE.json's stated rule).
 * ties_out (nullable, [B]) counts rounds in which the maximum was attained by more than one point, so a parity
 * test can report where the two rules could differ instead of hiding it.
 */
void oracle_fps_lowidx(const float *xyz, int pt_stride, int B, int N, int M, int32_t *idx_out,
                       int32_t *ties_out) {
    if (M <= 0) return;
#pragma omp parallel for schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        const float *p = xyz + (size_t)b * N * pt_stride;
        int32_t *out = idx_out + (size_t)b * M;
        float *temp = (float *)malloc(sizeof(float) * (size_t)(N > 0 ? N : 1));
        for (int k = 0; k < N; ++k) temp[k] = 1e10f;
        int old = 0, ties = 0;
        out[0] = old;
        for (int j = 1; j < M; ++j) {
            float x1 = p[(size_t)old * pt_stride + 0], y1 = p[(size_t)old * pt_stride + 1],
                  z1 = p[(size_t)old * pt_stride + 2];
            int besti = 0, nbest = 0;
            float best = -1.0f;
            for (int k = 0; k < N; ++k) {
                const float *q = p + (size_t)k * pt_stride;
                float d = sqdist_pn2(q[0] - x1, q[1] - y1, q[2] - z1);
                float d2 = d < temp[k] ? d : temp[k];
                temp[k] = d2;
                if (d2 > best) { best = d2; besti = k; nbest = 1; }
                else if (d2 == best) ++nbest;
            }
            if (nbest > 1) ++ties;
            old = besti;
            out[j] = old;
        }
        if (ties_out) ties_out[b] = ties;
        free(temp);
    }
}

/* out[b,c,j] = points[b,c,idx[b,j]]  (sampling_gpu.cu:15-23) */
void oracle_gather(const float *points, const int32_t *idx, int B, int C, int N, int M, float *out) {
    for (int b = 0; b < B; ++b)
        for (int c = 0; c < C; ++c)
            for (int j = 0; j < M; ++j)
                out[((size_t)b * C + c) * M + j] = points[((size_t)b * C + c) * N + idx[(size_t)b * M + j]];
}

/*
 * Ball query (ball_query_gpu.cu:23-44).  idx_out [B,M,ns] is zero-filled first like the caller does
 * (pointnet2_utils.py:218).  cnt_out (nullable, [B,M]) = number of genuine hits written (<= ns).
 */
void oracle_ball_query(const float *new_xyz, const float *xyz, int B, int N, int M, float radius, int ns,
                       int32_t *idx_out, int32_t *cnt_out) {
    float radius2 = radius * radius;
    memset(idx_out, 0, sizeof(int32_t) * (size_t)B * M * ns);
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; ++b) {
        for (int i = 0; i < M; ++i) {
            const float *c = new_xyz + ((size_t)b * M + i) * 3;
            const float *p = xyz + (size_t)b * N * 3;
            int32_t *idx = idx_out + ((size_t)b * M + i) * ns;
            float nx = c[0], ny = c[1], nz = c[2];
            int cnt = 0;
            for (int k = 0; k < N; ++k) {
                float d2 = sqdist_pn2(nx - p[k * 3 + 0], ny - p[k * 3 + 1], nz - p[k * 3 + 2]);
                if (d2 < radius2) {
                    if (cnt == 0)
                        for (int l = 0; l < ns; ++l) idx[l] = k;
                    idx[cnt] = k;
                    ++cnt;
                    if (cnt >= ns) break;
                }
            }
            if (cnt_out) cnt_out[(size_t)b * M + i] = cnt;
        }
    }
}

/* out[b,c,i,j] = points[b,c,idx[b,i,j]]  (group_points_gpu.cu:53-65) */
void oracle_group(const float *points, const int32_t *idx, int B, int C, int N, int M, int ns, float *out) {
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; ++b)
        for (int c = 0; c < C; ++c) {
            const float *src = points + ((size_t)b * C + c) * N;
            for (int i = 0; i < M; ++i)
                for (int j = 0; j < ns; ++j)
                    out[(((size_t)b * C + c) * M + i) * ns + j] = src[idx[((size_t)b * M + i) * ns + j]];
        }
}

/*
 * torch-cluster 1.5.9 knn(x, y, k, batch_x, batch_y) for equal-sized sorted batches
 * (deepclr.py:150-166): for every y row (query) the k nearest x rows of the same batch element.
 * x: [B*Nx,3], y: [B*Ny,3]; row_out/col_out: [B*Ny*k] int64, row = global y row id, col = global x row id,
 * ascending distance per query, ties -> lower x index (strict '>' insertion while scanning x ascending).
 * Requires Nx >= k (the reference's .view at deepclr.py:167 needs exactly k results per query).
 */
int oracle_knn(const float *x, const float *y, int B, int Nx, int Ny, int k, int64_t *row_out,
               int64_t *col_out) {
    if (k > Nx || k <= 0) return -1;
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; ++b) {
        for (int q = 0; q < Ny; ++q) {
            size_t gy = (size_t)b * Ny + q;
            float *dist = (float *)malloc(sizeof(float) * (size_t)k);
            int64_t *col = col_out + gy * k;
            for (int j = 0; j < k; ++j) { dist[j] = 1e38f; col[j] = -1; row_out[gy * k + j] = (int64_t)gy; }
            float qx = y[gy * 3 + 0], qy = y[gy * 3 + 1], qz = y[gy * 3 + 2];
            for (int c = 0; c < Nx; ++c) {
                size_t gx = (size_t)b * Nx + c;
                float dx = x[gx * 3 + 0] - qx, dy = x[gx * 3 + 1] - qy, dz = x[gx * 3 + 2] - qz;
                float d = dx * dx;       /* tmp_dist = 0 + dx*dx */
                d = fmaf(dy, dy, d);
                d = fmaf(dz, dz, d);
                for (int a = 0; a < k; ++a) {
                    if (dist[a] > d) {
                        for (int s = k - 1; s > a; --s) { dist[s] = dist[s - 1]; col[s] = col[s - 1]; }
                        dist[a] = d;
                        col[a] = (int64_t)gx;
                        break;
                    }
                }
            }
            free(dist);
        }
    }
    return 0;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
