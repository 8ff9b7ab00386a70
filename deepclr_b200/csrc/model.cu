// Model handle + forward orchestration: the C-ABI face of DeepCLR.forward / cloud_features
// (reference deepclr/models/deepclr.py:488-521).  One call enqueues the whole chain on the caller's stream.
#include "common.cuh"
#include "kernels.cuh"
#include "mlp_tc.cuh"
#include <nvtx3/nvToolsExt.h>
#include <new>
#include <stdlib.h>
#include <vector>

struct dclr_model {
    dclr_arch arch;
    int label_dim, F, nmerge_in, head_in;
    float *d_w = nullptr;            // all parameters, state-dict order
    size_t n_w = 0;
    std::vector<float> sa_w_host[2];   // SA MLP parameters, host copy (passed as kernel parameters)
    uint16_t *sa_tc_w = nullptr;       // the same as packed fp16 hi/lo images for sa_tc_kernel (sa_tc.cu)
    unsigned long long *sa_rows = nullptr;   // device counter of the density probe
    int sa_dense = -1, sa_dense_G = 0, sa_dense_N = 0;   // -1 unknown | 0 sparse: sa_mlp_kernel | 1 dense: sa_tc_kernel, for (G, N)
    struct Lin { const float *W, *b; int out, in; };
    Lin merge[3], outc[5], lin[2], fin;
    int numerics = 0;
    int tie_rule = DCLR_TIE_REFERENCE;   // FPS tie order (dclr_model_set_fps_tie_rule)
    // workspace
    void *ws = nullptr;
    size_t ws_bytes = 0;
    int ws_generation = 0;           // bumped whenever the workspace moves: graphs captured before are stale
    int last_launches = 0;
    // carved pointers (valid after ensure_ws)
    int32_t *fps_idx, *bq_idx[2], *bq_cnt[2], *knn_col;
    float4 *knn_nbr4;          // [B,M,k] {source - template displacement, source index}: the flow kernel's row records
    void *bq_ws, *fps_ws;
    float *new_xyz, *sa_feat, *fps_temp, *rows, *h1, *h2, *h3, *motion, *actA, *actB, *pooled, *t1, *t2;
    uint8_t *mask;
    uint16_t *sa_feat16, *motion16;  // 16-bit images of sa_feat / motion for the tensor-core stages
    dclr::TcModel *tc = nullptr;     // bf16 tensor-core path (mlp_tc.cu)
    // optional per-stage CUDA-event timing (dclr_model_set_timing): ring of calls x (stages + 1) events
    bool timing = false;
    cudaEvent_t *ev = nullptr;
    // side stream: work that reads only the input cloud runs next to FPS (which leaves 20 of the 148 SMs idle)
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_fps = nullptr, ev_knn = nullptr;
    int ev_calls = 0;                // calls recorded since the last collect (may exceed the ring)
    int ev_cur = 0;                  // slot of the call being recorded
    uint32_t ev_mask[64] = {0};      // stages recorded per ring slot
};

namespace dclr {

constexpr int SA_TC_MIN_ROWS_PER_CENTRE = 8;   // mean genuine rows per (centre, scale) from which sa_tc_kernel is used

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int label_dim_of(int t) { return t == 0 ? 6 : (t == 1 ? 7 : 8); }

// carve (or just measure when base == nullptr) the workspace for G clouds of N points / B pairs
static size_t carve(dclr_model *m, char *base, int G, int N, int B) {
    const dclr_arch &a = m->arch;
    const int M = a.npoint;
    size_t off = 0;
    auto take = [&](size_t bytes) { char *p = base ? base + off : nullptr; off = align_up(off + bytes, 256); return p; };
    m->fps_idx = (int32_t *)take((size_t)G * M * 4);
    m->new_xyz = (float *)take((size_t)G * M * 3 * 4);
    for (int s = 0; s < 2; ++s) {
        int ns = s < a.num_scales ? a.nsamples[s] : 0;
        m->bq_idx[s] = (int32_t *)take((size_t)G * M * ns * 4);
        m->bq_cnt[s] = (int32_t *)take((size_t)G * M * 4);
    }
    m->sa_feat = (float *)take((size_t)G * M * m->F * 4);
    m->fps_temp = (float *)take(N > 16384 ? (size_t)G * N * 4 : 0);
    m->bq_ws = take(ball_grid_workspace_bytes(G, N));
    m->fps_ws = take(fps_bucket_workspace_bytes(G, N));
    const size_t BM = (size_t)B * M;
    const size_t R = BM * (a.knn_k > 0 ? a.knn_k : 1);
    m->knn_col = (int32_t *)take(R * 4);
    m->motion = (float *)take(BM * align_up(3 + a.merge_mlp[2], 4) * 4);
    m->pooled = (float *)take((size_t)B * a.out_mlp[4] * 4);
    m->t1 = (float *)take((size_t)B * a.out_linear[1] * 4);
    m->t2 = (float *)take((size_t)B * a.out_linear[2] * 4);
    if (m->numerics == 0) {
        int wmax = 0;
        for (int i = 0; i < 5; ++i) wmax = a.out_mlp[i] > wmax ? a.out_mlp[i] : wmax;
        m->rows = (float *)take(R * align_up(m->nmerge_in, 4) * 4);
        m->h1 = (float *)take(R * a.merge_mlp[0] * 4);
        m->h2 = (float *)take(R * a.merge_mlp[1] * 4);
        m->h3 = (float *)take(R * a.merge_mlp[2] * 4);
        m->mask = (uint8_t *)take(R);
        m->actA = (float *)take(BM * wmax * 4);
        m->actB = (float *)take(BM * wmax * 4);
        m->sa_feat16 = m->motion16 = nullptr;
        m->knn_nbr4 = nullptr;
    } else {
        m->rows = m->h1 = m->h2 = m->h3 = m->actA = m->actB = nullptr;
        m->mask = nullptr;
        m->sa_feat16 = (uint16_t *)take((size_t)G * M * m->F * 2);
        m->motion16 = (uint16_t *)take(BM * a.merge_mlp[2] * 2);
        m->knn_nbr4 = (float4 *)take(R * 16);
    }
    return off;
}

static int ensure_ws(dclr_model *m, int G, int N, int B) {
    size_t need = carve(m, nullptr, G, N, B);
    if (need > m->ws_bytes) {
        if (m->ws) DCLR_CUDA(cudaFree(m->ws));
        m->ws = nullptr;
        m->ws_bytes = 0;
        DCLR_CUDA(cudaMalloc(&m->ws, need));
        m->ws_bytes = need;
        ++m->ws_generation;
    }
    carve(m, (char *)m->ws, G, N, B);
    return DCLR_OK;
}

// stages a forward is split into for timing
enum { ST_FPS = 0, ST_BALL, ST_SA, ST_KNN, ST_FLOW, ST_HEAD, ST_TAIL, ST_LAYOUT, ST_COUNT };
static const char *const kStageNames[ST_COUNT] = {"fps", "ball_query", "sa_mlp_maxpool", "knn",
                                                  "flow_embed", "pose_head_conv", "pose_tail", "layout"};
constexpr int EV_RING = 64;
constexpr int EV_PER_CALL = 2 * ST_COUNT;      // begin/end per stage

static void timing_begin_call(dclr_model *m) {
    if (!m->timing) return;
    m->ev_cur = m->ev_calls % EV_RING;
    ++m->ev_calls;
    m->ev_mask[m->ev_cur] = 0;
}
// records begin (which = 0) / end (which = 1) of stage s; a stage hit twice in one call keeps the first begin/last end
static void timing_mark(dclr_model *m, int s, int which, cudaStream_t st) {
    if (!m->timing) return;
    if (which == 0 && (m->ev_mask[m->ev_cur] >> s & 1u)) return;
    m->ev_mask[m->ev_cur] |= 1u << s;
    cudaEventRecord(m->ev[(size_t)m->ev_cur * EV_PER_CALL + 2 * s + which], st);
}
// one stage of a forward: CUDA-event pair when timing is on (the reference times with an event pair around predict,
// scripts/timing.py:32-47) + an NVTX range named after the stage for timeline tools (free when no tool is attached)
struct StageScope {
    dclr_model *m; int s; cudaStream_t st;
    StageScope(dclr_model *m_, int s_, cudaStream_t st_) : m(m_), s(s_), st(st_) {
        nvtxRangePushA(kStageNames[s]);
        timing_mark(m, s, 0, st);
    }
    ~StageScope() {
        timing_mark(m, s, 1, st);
        nvtxRangePop();
    }
};

#define DCLR_TRY(expr)            \
    do {                          \
        int rc__ = (expr);        \
        if (rc__ != DCLR_OK) return rc__; \
        ++m->last_launches;       \
    } while (0)

// FPS -> ball query (both radii, one pass) -> grouped MLP + max-pool.  pointnet2_modules.py:19-55
// knn_pairs > 0: the kNN of the merge stage (it reads only the FPS samples) is forked onto the side stream right after
// FPS, next to ball query + grouped MLP; the caller joins with ev_knn before the flow embedding.
// centres != nullptr: the caller supplies the sampled centres [G,npoint,3] (dclr_sa_msg_fused): no FPS.
static int run_set_abstraction(dclr_model *m, const float *x, int G, int N, cudaStream_t st, int knn_pairs = 0,
                               const float *centres = nullptr) {
    const dclr_arch &a = m->arch;
    // The ball-query grid build (counting sort of the cloud into cells, ~40 us for 128 KITTI clouds) depends on the
    // input only, and the sampling kernel (one CTA per cloud, latency-bound) leaves SMs and issue slots idle for
    // ~0.5 ms: fork the build onto a low-priority side stream once the FPS sort is done, join before the query.
    if (!m->side) {
        int lo = 0, hi = 0;
        DCLR_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));     // lo = least priority
        DCLR_CUDA(cudaStreamCreateWithPriority(&m->side, cudaStreamNonBlocking, lo));
        DCLR_CUDA(cudaEventCreateWithFlags(&m->ev_fork, cudaEventDisableTiming));
        DCLR_CUDA(cudaEventCreateWithFlags(&m->ev_join, cudaEventDisableTiming));
        DCLR_CUDA(cudaEventCreateWithFlags(&m->ev_fps, cudaEventDisableTiming));
        DCLR_CUDA(cudaEventCreateWithFlags(&m->ev_knn, cudaEventDisableTiming));
    }
    if (centres) {
        DCLR_CUDA(cudaMemcpyAsync(m->new_xyz, centres, (size_t)G * a.npoint * 3 * sizeof(float), cudaMemcpyDeviceToDevice, st));
        DCLR_CUDA(cudaEventRecord(m->ev_fork, st));
    } else {
        StageScope sc(m, ST_FPS, st);
        int used = 0;
        DCLR_TRY(fps_bucket_launch(x, a.input_dim, G, N, a.npoint, m->tie_rule, m->fps_idx, m->new_xyz, m->fps_ws, st, &used,
                                   m->ev_fork));
        if (used) ++m->last_launches;                      // sort + sample
        else {
            DCLR_CUDA(cudaEventRecord(m->ev_fork, st));
            DCLR_TRY(fps_launch(x, a.input_dim, G, N, a.npoint, m->tie_rule, m->fps_idx, m->new_xyz, m->fps_temp, st));
        }
    }
    {
        int grid_built = 0;
        DCLR_CUDA(cudaStreamWaitEvent(m->side, m->ev_fork, 0));
        DCLR_TRY(ball_grid_launch(nullptr, x, a.input_dim, G, N, a.npoint, a.num_scales, a.radii, a.nsamples, m->bq_idx,
                                  m->bq_cnt, /*pad=*/0, m->bq_ws, m->side, &grid_built, BG_PHASE_BUILD));
        if (!grid_built) --m->last_launches;               // shape not handled by the grid: nothing was launched
        DCLR_CUDA(cudaEventRecord(m->ev_join, m->side));
    }
    if (knn_pairs > 0) {
        // candidates x = source samples (clouds B..2B-1), queries y = template samples (deepclr.py:164-166)
        const size_t BM = (size_t)knn_pairs * a.npoint;
        DCLR_CUDA(cudaEventRecord(m->ev_fps, st));
        DCLR_CUDA(cudaStreamWaitEvent(m->side, m->ev_fps, 0));
        {
            StageScope sc(m, ST_KNN, m->side);
            DCLR_TRY(knn_launch(m->new_xyz + BM * 3, m->new_xyz, knn_pairs, a.npoint, a.npoint, a.knn_k, nullptr, nullptr,
                                m->knn_col, m->knn_nbr4, m->side));
        }
        DCLR_CUDA(cudaEventRecord(m->ev_knn, m->side));
    }
    {
        DCLR_CUDA(cudaStreamWaitEvent(st, m->ev_join, 0));
        StageScope sc(m, ST_BALL, st);
        int used = 0;
        DCLR_TRY(ball_grid_launch(m->new_xyz, x, a.input_dim, G, N, a.npoint, a.num_scales, a.radii, a.nsamples,
                                  m->bq_idx, m->bq_cnt, /*pad=*/0, m->bq_ws, st, &used, BG_PHASE_QUERY));
        if (!used) --m->last_launches;
        if (!used)
            DCLR_TRY(ball_query_launch(m->new_xyz, x, a.input_dim, G, N, a.npoint, a.num_scales, a.radii, a.nsamples,
                                       m->bq_idx, m->bq_cnt, /*pad=*/0, st));
    }
    StageScope sc(m, ST_SA, st);
    SaArgs s;
    s.x = x; s.new_xyz = m->new_xyz; s.N = N; s.M = a.npoint; s.ld_out = m->F; s.nscales = a.num_scales;
    s.out = m->sa_feat;
    s.out16 = m->numerics != 0 ? m->sa_feat16 : nullptr;
    s.fmt16 = m->numerics;
    for (int i = 0; i < 2; ++i) {
        s.idx[i] = m->bq_idx[i]; s.cnt[i] = m->bq_cnt[i]; s.ns[i] = a.nsamples[i < a.num_scales ? i : 0];
        s.w_host[i] = m->sa_w_host[i < a.num_scales ? i : 0].data();
    }
    // Which kernel?  fp32 numerics: always the fp32 SIMT kernel.  16-bit modes: the tensor-core kernel (fp16 hi/lo split,
    // fp32-class results) when the neighbourhoods are dense -- measured once per (clouds, points) shape of a handle: the
    // first call counts the genuine rows (bq_cnt), reads the total back (one synchronisation, never during a stream
    // capture) and keeps the decision.  Crossover: tools/sa_density_sweep.py.  DCLR_SA_ROUTE=simt|tc overrides.
    int route = 0;
    if (m->numerics != 0) {
        const char *env = getenv("DCLR_SA_ROUTE");
        if (env && env[0] == 't') route = 1;
        else if (env && env[0] == 's') route = 0;
        else if ((long long)G * a.npoint < 148LL * 128) {
            // small launches (up to 9 KITTI pairs): latency matters, not what shares the SMs with the sampling kernel of
            // another step, and the tensor kernel is the shorter one at every density (batch 1: 0.030 -> 0.019 ms)
            route = 1;
        } else {
            if (m->sa_dense < 0 || m->sa_dense_G != G || m->sa_dense_N != N) {
                cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
                DCLR_CUDA(cudaStreamIsCapturing(st, &cap));
                if (cap == cudaStreamCaptureStatusNone) {
                    unsigned long long rows = 0;
                    DCLR_CUDA(cudaMemsetAsync(m->sa_rows, 0, sizeof(rows), st));
                    for (int i = 0; i < a.num_scales; ++i)
                        DCLR_TRY(sa_count_rows_launch(m->bq_cnt[i], (size_t)G * a.npoint, m->sa_rows, st));
                    DCLR_CUDA(cudaMemcpyAsync(&rows, m->sa_rows, sizeof(rows), cudaMemcpyDeviceToHost, st));
                    DCLR_CUDA(cudaStreamSynchronize(st));
                    m->sa_dense = rows >= (unsigned long long)SA_TC_MIN_ROWS_PER_CENTRE * G * a.npoint * a.num_scales;
                    m->sa_dense_G = G; m->sa_dense_N = N;
                    m->last_launches -= a.num_scales;              // (the probe is not part of a steady-state forward)
                }
            }
            route = m->sa_dense > 0;
        }
    }
    if (route) DCLR_TRY(sa_tc_launch(s, m->sa_tc_w, G, a.input_dim, st));
    else DCLR_TRY(sa_mlp_launch(s, G, a.input_dim, a.sa_mlp[0], st));
    return DCLR_OK;
}

// kNN grouping + motion embedding on point-major SA features (deepclr.py:157-231): leaves the motion features in the
// workspace (16-bit image for the tensor-core head, fp32 `motion` when tapped / in fp32 mode)
static int run_flow(dclr_model *m, int B, bool tap_motion, cudaStream_t st, bool knn_forked) {
    const dclr_arch &a = m->arch;
    const int M = a.npoint, k = a.knn_k, F = m->F;
    const size_t BM = (size_t)B * M, R = BM * k;
    const int ldm = (int)align_up(3 + a.merge_mlp[2], 4);
    if (knn_forked) {
        DCLR_CUDA(cudaStreamWaitEvent(st, m->ev_knn, 0));
    } else {
        // candidates x = source samples (clouds B..2B-1), queries y = template samples (deepclr.py:164-166)
        StageScope sc(m, ST_KNN, st);
        DCLR_TRY(knn_launch(m->new_xyz + BM * 3, m->new_xyz, B, M, M, k, nullptr, nullptr, m->knn_col, m->knn_nbr4, st));
    }
    StageScope sc(m, ST_FLOW, st);
    if (m->numerics != 0) {
        // fp32 motion features only when the caller taps them; the head reads the 16-bit image
        DCLR_TRY(tc_flow(m->tc, m->sa_feat16, m->knn_nbr4, B, M, m->motion16, tap_motion ? m->motion : nullptr, st));
        return DCLR_OK;
    }
    const int ld0 = (int)align_up(m->nmerge_in, 4);
    DCLR_TRY(flow_rows_launch(m->new_xyz, m->sa_feat, m->knn_col, B, M, k, F, a.merge_radius, m->rows, ld0, m->mask, st));
    DCLR_TRY(linear_f32_launch(m->rows, ld0, m->merge[0].W, m->merge[0].b, m->h1, a.merge_mlp[0], (int)R, a.merge_mlp[0],
                               m->nmerge_in, 1, st));
    DCLR_TRY(linear_f32_launch(m->h1, a.merge_mlp[0], m->merge[1].W, m->merge[1].b, m->h2, a.merge_mlp[1], (int)R,
                               a.merge_mlp[1], a.merge_mlp[0], 1, st));
    DCLR_TRY(linear_f32_launch(m->h2, a.merge_mlp[1], m->merge[2].W, m->merge[2].b, m->h3, a.merge_mlp[2], (int)R,
                               a.merge_mlp[2], a.merge_mlp[1], 1, st));
    DCLR_TRY(masked_groupmax_launch(m->h3, a.merge_mlp[2], m->mask, m->new_xyz, BM, k, a.merge_mlp[2], m->motion, ldm, st));
    return DCLR_OK;
}

// reference layout [B, 3 + merge_mlp[2], npoint] of the motion features left by run_flow
static int motion_to_reference_layout(dclr_model *m, int B, float *motion_out, cudaStream_t st) {
    const dclr_arch &a = m->arch;
    StageScope sc(m, ST_LAYOUT, st);
    if (m->numerics != 0)    // tensor-core path keeps features only; coordinates are the template samples
        DCLR_TRY(pm_to_cm_launch(m->new_xyz, m->motion, a.merge_mlp[2], 0, B, a.npoint, a.merge_mlp[2], motion_out, st));
    else
        DCLR_TRY(pm_to_cm_launch(nullptr, m->motion, (int)align_up(3 + a.merge_mlp[2], 4), 0, B, a.npoint, a.merge_mlp[2],
                                 motion_out, st));
    return DCLR_OK;
}

// output head on the motion features in the workspace (deepclr.py:284-294): 5 x (1x1 conv + ReLU) -> max over the points
// -> 2 x (Linear + ReLU) -> Linear -> activation
static int run_head(dclr_model *m, int B, float *y_out, cudaStream_t st) {
    const dclr_arch &a = m->arch;
    const int M = a.npoint;
    const size_t BM = (size_t)B * M;
    const int ldm = (int)align_up(3 + a.merge_mlp[2], 4);
    if (m->numerics != 0) {
        StageScope sc(m, ST_HEAD, st);
        DCLR_TRY(tc_head(m->tc, m->new_xyz, m->motion16, B, M, m->pooled, st));
    } else {
        StageScope sc(m, ST_HEAD, st);
        const float *in = m->motion;
        int ldin = ldm, kin = m->head_in;
        float *bufs[2] = {m->actA, m->actB};
        for (int l = 0; l < 5; ++l) {
            float *out = bufs[l & 1];
            DCLR_TRY(linear_f32_launch(in, ldin, m->outc[l].W, m->outc[l].b, out, a.out_mlp[l], (int)BM, a.out_mlp[l],
                                       kin, 1, st));
            in = out; ldin = a.out_mlp[l]; kin = a.out_mlp[l];
        }
        DCLR_TRY(rowmax_launch(in, ldin, B, M, a.out_mlp[4], m->pooled, st));
    }
    StageScope sc(m, ST_TAIL, st);
    // fully connected tail: helper.py:97-123 (ReLU after both), final nn.Linear, activation deepclr.py:275-282
    if (a.out_linear[0] % 4 == 0 && a.out_linear[1] % 4 == 0 && a.out_linear[2] % 4 == 0) {
        DCLR_TRY(linear_skinny_launch(m->pooled, a.out_mlp[4], m->lin[0].W, m->lin[0].b, m->t1, a.out_linear[1], B,
                                      a.out_linear[1], a.out_linear[0], 1, 0, st));
        DCLR_TRY(linear_skinny_launch(m->t1, a.out_linear[1], m->lin[1].W, m->lin[1].b, m->t2, a.out_linear[2], B,
                                      a.out_linear[2], a.out_linear[1], 1, 0, st));
        DCLR_TRY(linear_skinny_launch(m->t2, a.out_linear[2], m->fin.W, m->fin.b, y_out, m->label_dim, B,
                                      m->label_dim, a.out_linear[2], 2, a.label_type, st));
    } else {
        DCLR_TRY(linear_f32_launch(m->pooled, a.out_mlp[4], m->lin[0].W, m->lin[0].b, m->t1, a.out_linear[1], B,
                                   a.out_linear[1], a.out_linear[0], 1, st));
        DCLR_TRY(linear_f32_launch(m->t1, a.out_linear[1], m->lin[1].W, m->lin[1].b, m->t2, a.out_linear[2], B,
                                   a.out_linear[2], a.out_linear[1], 1, st));
        DCLR_TRY(linear_f32_launch(m->t2, a.out_linear[2], m->fin.W, m->fin.b, y_out, m->label_dim, B, m->label_dim,
                                   a.out_linear[2], 0, st));
        DCLR_TRY(pose_activation_launch(y_out, B, m->label_dim, a.label_type, st));
    }
    return DCLR_OK;
}

static int run_merge_output(dclr_model *m, int B, float *y_out, float *motion_out, cudaStream_t st,
                            bool knn_forked = false) {
    int rc = run_flow(m, B, motion_out != nullptr, st, knn_forked);
    if (rc != DCLR_OK) return rc;
    rc = run_head(m, B, y_out, st);
    if (rc != DCLR_OK) return rc;
    return motion_out ? motion_to_reference_layout(m, B, motion_out, st) : DCLR_OK;
}

}  // namespace dclr

// ------------------------------------------------------------------ C ABI
using namespace dclr;

extern "C" int dclr_model_create(const dclr_arch *arch, const float *weights, size_t num_floats, dclr_model **out) {
    DCLR_CHECK_ARG(arch && weights && out, "dclr_model_create: null pointer");
    const dclr_arch &a = *arch;
    DCLR_CHECK_ARG(a.input_dim == 3 || a.input_dim == 4, "input_dim must be 3 or 4 (got %d)", a.input_dim);
    DCLR_CHECK_ARG(a.num_scales == 1 || a.num_scales == 2, "num_scales must be 1 or 2");
    DCLR_CHECK_ARG(a.npoint >= 1 && a.knn_k >= 1 && a.knn_k <= 128, "need npoint >= 1 and 1 <= k <= 128");
    DCLR_CHECK_ARG(a.label_type >= 0 && a.label_type <= 2, "label_type must be 0..2");
    DCLR_CHECK_ARG(a.out_linear[0] == a.out_mlp[4], "out_linear[0] must equal out_mlp[4]");
    for (int s = 0; s < a.num_scales; ++s)       // the grouped-MLP kernel is built for the reference widths only (sa_mlp.cu)
        if (a.sa_mlp[s][0] != 16 || a.sa_mlp[s][1] != 16 || a.sa_mlp[s][2] != 32)
            return fail(DCLR_E_UNSUPPORTED, "dclr_model_create: set-abstraction MLP of scale %d is [%d,%d,%d]; only "
                        "[16,16,32] is built", s, a.sa_mlp[s][0], a.sa_mlp[s][1], a.sa_mlp[s][2]);
    dclr_model *m = new (std::nothrow) dclr_model();
    DCLR_CHECK_ARG(m, "out of host memory");
    m->arch = a;
    m->label_dim = label_dim_of(a.label_type);
    m->F = 0;
    for (int s = 0; s < a.num_scales; ++s) m->F += a.sa_mlp[s][2];
    m->nmerge_in = 3 + 2 * m->F;                 // append_features=True (deepclr.py:193-194)
    m->head_in = 3 + a.merge_mlp[2];
    // walk the state-dict order and record offsets
    size_t off = 0;
    std::vector<size_t> w_off, b_off;
    auto add = [&](int o, int i) { w_off.push_back(off); off += (size_t)o * i; b_off.push_back(off); off += o; };
    size_t sa_off[2] = {0, 0};
    for (int s = 0; s < a.num_scales; ++s) {
        sa_off[s] = off;
        int cin = a.input_dim;                   // (input_dim - 3) features + 3 xyz (pointnet2_modules.py:88-89)
        for (int l = 0; l < 3; ++l) { add(a.sa_mlp[s][l], cin); cin = a.sa_mlp[s][l]; }
    }
    size_t first_dense = w_off.size();
    int cin = m->nmerge_in;
    for (int l = 0; l < 3; ++l) { add(a.merge_mlp[l], cin); cin = a.merge_mlp[l]; }
    cin = m->head_in;
    for (int l = 0; l < 5; ++l) { add(a.out_mlp[l], cin); cin = a.out_mlp[l]; }
    for (int l = 0; l < 2; ++l) add(a.out_linear[l + 1], a.out_linear[l]);
    add(m->label_dim, a.out_linear[2]);
    if (off != num_floats) {
        delete m;
        return fail(DCLR_E_INVALID, "dclr_model_create: expected %zu parameters for this architecture, got %zu", off,
                    num_floats);
    }
    cudaError_t e = cudaMalloc(&m->d_w, off * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(m->d_w, weights, off * sizeof(float), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        if (m->d_w) cudaFree(m->d_w);
        delete m;
        return fail(DCLR_E_CUDA, "dclr_model_create: %s", cudaGetErrorString(e));
    }
    m->n_w = off;
    for (int s = 0; s < a.num_scales; ++s) {
        const size_t end = s + 1 < a.num_scales ? sa_off[s + 1] : w_off[first_dense];
        m->sa_w_host[s].assign(weights + sa_off[s], weights + end);
    }
    {
        // tensor-core images of the SA weights (used in the 16-bit numerics modes on dense neighbourhoods)
        std::vector<uint16_t> img(sa_tc_weight_bytes() / sizeof(uint16_t), 0);
        for (int s = 0; s < a.num_scales; ++s) sa_tc_pack_scale(m->sa_w_host[s].data(), a.input_dim, img.data() + s * (img.size() / 2));
        e = cudaMalloc(&m->sa_tc_w, sa_tc_weight_bytes());
        if (e == cudaSuccess) e = cudaMemcpy(m->sa_tc_w, img.data(), sa_tc_weight_bytes(), cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = cudaMalloc(&m->sa_rows, sizeof(unsigned long long));
        if (e != cudaSuccess) {
            dclr_model_destroy(m);
            return fail(DCLR_E_CUDA, "dclr_model_create: %s", cudaGetErrorString(e));
        }
    }
    size_t q = first_dense;
    auto lin = [&](int o, int i) { dclr_model::Lin L{m->d_w + w_off[q], m->d_w + b_off[q], o, i}; ++q; return L; };
    cin = m->nmerge_in;
    for (int l = 0; l < 3; ++l) { m->merge[l] = lin(a.merge_mlp[l], cin); cin = a.merge_mlp[l]; }
    cin = m->head_in;
    for (int l = 0; l < 5; ++l) { m->outc[l] = lin(a.out_mlp[l], cin); cin = a.out_mlp[l]; }
    for (int l = 0; l < 2; ++l) m->lin[l] = lin(a.out_linear[l + 1], a.out_linear[l]);
    m->fin = lin(m->label_dim, a.out_linear[2]);
    *out = m;
    return DCLR_OK;
}

extern "C" void dclr_model_destroy(dclr_model *m) {
    if (!m) return;
    if (m->tc) tc_destroy(m->tc);
    if (m->ev) {
        for (int i = 0; i < EV_RING * EV_PER_CALL; ++i) cudaEventDestroy(m->ev[i]);
        delete[] m->ev;
    }
    if (m->ws) cudaFree(m->ws);
    if (m->d_w) cudaFree(m->d_w);
    if (m->sa_tc_w) cudaFree(m->sa_tc_w);
    if (m->sa_rows) cudaFree(m->sa_rows);
    if (m->ev_fork) cudaEventDestroy(m->ev_fork);
    if (m->ev_join) cudaEventDestroy(m->ev_join);
    if (m->ev_fps) cudaEventDestroy(m->ev_fps);
    if (m->ev_knn) cudaEventDestroy(m->ev_knn);
    if (m->side) cudaStreamDestroy(m->side);
    delete m;
}

extern "C" int dclr_model_set_numerics(dclr_model *m, int mode) {
    DCLR_CHECK_ARG(m, "null model");
    DCLR_CHECK_ARG(mode >= 0 && mode <= 2, "numerics mode must be 0 (fp32), 1 (bf16) or 2 (fp16 tensor cores)");
    if (mode != 0 && m->tc && tc_format(m->tc) != mode) {
        tc_destroy(m->tc);
        m->tc = nullptr;
    }
    if (mode != 0 && !m->tc) {
        TcLayerSpec specs[8];
        for (int l = 0; l < 3; ++l) specs[l] = {m->merge[l].W, m->merge[l].b, m->merge[l].out, m->merge[l].in};
        for (int l = 0; l < 5; ++l) specs[3 + l] = {m->outc[l].W, m->outc[l].b, m->outc[l].out, m->outc[l].in};
        int rc = tc_create(specs, m->F, m->arch.knn_k, m->arch.merge_radius, mode, &m->tc);
        if (rc != DCLR_OK) return rc;
    }
    m->numerics = mode;
    return DCLR_OK;
}

extern "C" int dclr_model_set_fps_tie_rule(dclr_model *m, int rule) {
    DCLR_CHECK_ARG(m, "null model");
    DCLR_CHECK_ARG(rule == DCLR_TIE_LOWEST_INDEX || rule == DCLR_TIE_REFERENCE,
                   "dclr_model_set_fps_tie_rule: rule must be 0 (lowest index) or 1 (reference kernel order), got %d", rule);
    m->tie_rule = rule;
    return DCLR_OK;
}
extern "C" int dclr_model_get_fps_tie_rule(const dclr_model *m) { return m ? m->tie_rule : -1; }

extern "C" int dclr_cloud_features(dclr_model *m, const float *x, int num_clouds, int N, float *feat_out,
                                   void *stream) {
    DCLR_CHECK_ARG(m && x && feat_out, "dclr_cloud_features: null pointer");
    DCLR_CHECK_ARG(num_clouds >= 1 && N >= 1, "dclr_cloud_features: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    m->last_launches = 0;
    timing_begin_call(m);
    int rc = ensure_ws(m, num_clouds, N, (num_clouds + 1) / 2);
    if (rc != DCLR_OK) return rc;
    rc = run_set_abstraction(m, x, num_clouds, N, st);
    if (rc != DCLR_OK) return rc;
    StageScope sc(m, ST_LAYOUT, st);
    DCLR_TRY(pm_to_cm_launch(m->new_xyz, m->sa_feat, m->F, 0, num_clouds, m->arch.npoint, m->F, feat_out, st));
    return DCLR_OK;
}

extern "C" int dclr_forward_features(dclr_model *m, const float *feat, int B, float *y_out, float *motion_out,
                                     void *stream) {
    DCLR_CHECK_ARG(m && feat && y_out, "dclr_forward_features: null pointer");
    DCLR_CHECK_ARG(B >= 1, "dclr_forward_features: bad batch");
    DCLR_CHECK_ARG(m->arch.knn_k <= m->arch.npoint, "knn k=%d exceeds npoint=%d", m->arch.knn_k, m->arch.npoint);
    cudaStream_t st = (cudaStream_t)stream;
    m->last_launches = 0;
    timing_begin_call(m);
    int rc = ensure_ws(m, 2 * B, 1, B);
    if (rc != DCLR_OK) return rc;
    {
        StageScope sc(m, ST_LAYOUT, st);
        DCLR_TRY(cm_to_pm_launch(feat, 2 * B, m->arch.npoint, m->F, m->new_xyz, m->sa_feat, m->F, st));
        if (m->numerics != 0)
            DCLR_TRY(tc_convert(m->tc, m->sa_feat, m->sa_feat16, (size_t)2 * B * m->arch.npoint * m->F, st));
    }
    return run_merge_output(m, B, y_out, motion_out, st);
}

extern "C" int dclr_forward_pairs(dclr_model *m, const float *x, int B, int N, float *y_out, float *feat_out,
                                  float *motion_out, void *stream) {
    DCLR_CHECK_ARG(m && x && y_out, "dclr_forward_pairs: null pointer");
    DCLR_CHECK_ARG(B >= 1 && N >= 1, "dclr_forward_pairs: bad shape");
    DCLR_CHECK_ARG(m->arch.knn_k <= m->arch.npoint, "knn k=%d exceeds npoint=%d", m->arch.knn_k, m->arch.npoint);
    cudaStream_t st = (cudaStream_t)stream;
    m->last_launches = 0;
    timing_begin_call(m);
    int rc = ensure_ws(m, 2 * B, N, B);
    if (rc != DCLR_OK) return rc;
    rc = run_set_abstraction(m, x, 2 * B, N, st, /*knn_pairs=*/B);
    if (rc != DCLR_OK) return rc;
    if (feat_out) {
        StageScope sc(m, ST_LAYOUT, st);
        DCLR_TRY(pm_to_cm_launch(m->new_xyz, m->sa_feat, m->F, 0, 2 * B, m->arch.npoint, m->F, feat_out, st));
    }
    return run_merge_output(m, B, y_out, motion_out, st, /*knn_forked=*/true);
}

// ---- stage-level entries: the three fused stacks one by one, for op-level parity and sweeps
extern "C" size_t dclr_workspace_bytes(const dclr_model *m, int num_clouds, int N) {
    if (!m || num_clouds < 1 || N < 1) return 0;
    dclr_model probe = *m;                       // carve() only writes the carved pointers of its argument
    probe.tc = nullptr;
    const size_t need = carve(&probe, nullptr, num_clouds, N, (num_clouds + 1) / 2);
    probe.ev = nullptr; probe.ws = nullptr; probe.d_w = nullptr; probe.side = nullptr;
    return need;
}

extern "C" int dclr_sa_msg_fused(dclr_model *m, const float *x, const float *new_xyz, int num_clouds, int N,
                                 float *feat_out, void *stream) {
    DCLR_CHECK_ARG(m && x && new_xyz && feat_out, "dclr_sa_msg_fused: null pointer");
    DCLR_CHECK_ARG(num_clouds >= 1 && N >= 1, "dclr_sa_msg_fused: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    m->last_launches = 0;
    timing_begin_call(m);
    int rc = ensure_ws(m, num_clouds, N, (num_clouds + 1) / 2);
    if (rc != DCLR_OK) return rc;
    rc = run_set_abstraction(m, x, num_clouds, N, st, 0, new_xyz);
    if (rc != DCLR_OK) return rc;
    StageScope sc(m, ST_LAYOUT, st);             // features only, [G, F, npoint] as PointnetSAModuleMSG returns them
    DCLR_TRY(pm_to_cm_launch(nullptr, m->sa_feat, m->F, 0, num_clouds, m->arch.npoint, m->F - 3, feat_out, st));
    return DCLR_OK;
}

extern "C" int dclr_flow_embed_fused(dclr_model *m, const float *feat, int B, float *motion_out, void *stream) {
    DCLR_CHECK_ARG(m && feat && motion_out, "dclr_flow_embed_fused: null pointer");
    DCLR_CHECK_ARG(B >= 1, "dclr_flow_embed_fused: bad batch");
    DCLR_CHECK_ARG(m->arch.knn_k <= m->arch.npoint, "knn k=%d exceeds npoint=%d", m->arch.knn_k, m->arch.npoint);
    cudaStream_t st = (cudaStream_t)stream;
    m->last_launches = 0;
    timing_begin_call(m);
    int rc = ensure_ws(m, 2 * B, 1, B);
    if (rc != DCLR_OK) return rc;
    {
        StageScope sc(m, ST_LAYOUT, st);
        DCLR_TRY(cm_to_pm_launch(feat, 2 * B, m->arch.npoint, m->F, m->new_xyz, m->sa_feat, m->F, st));
        if (m->numerics != 0)
            DCLR_TRY(tc_convert(m->tc, m->sa_feat, m->sa_feat16, (size_t)2 * B * m->arch.npoint * m->F, st));
    }
    rc = run_flow(m, B, true, st, false);
    if (rc != DCLR_OK) return rc;
    return motion_to_reference_layout(m, B, motion_out, st);
}

extern "C" int dclr_pose_head_fused(dclr_model *m, const float *motion, int B, float *y_out, void *stream) {
    DCLR_CHECK_ARG(m && motion && y_out, "dclr_pose_head_fused: null pointer");
    DCLR_CHECK_ARG(B >= 1, "dclr_pose_head_fused: bad batch");
    cudaStream_t st = (cudaStream_t)stream;
    const dclr_arch &a = m->arch;
    m->last_launches = 0;
    timing_begin_call(m);
    int rc = ensure_ws(m, 2 * B, 1, B);
    if (rc != DCLR_OK) return rc;
    {
        StageScope sc(m, ST_LAYOUT, st);
        const size_t BM = (size_t)B * a.npoint;
        if (m->numerics != 0) {      // coordinates -> new_xyz, features -> fp32 rows -> the head's 16-bit operand image
            DCLR_TRY(cm_to_pm_launch(motion, B, a.npoint, a.merge_mlp[2], m->new_xyz, m->motion, a.merge_mlp[2], st));
            DCLR_TRY(tc_convert(m->tc, m->motion, m->motion16, BM * a.merge_mlp[2], st));
        } else {                     // rows [xyz | features] with the SIMT head's leading dimension
            DCLR_TRY(cm_to_pm_launch(motion, B, a.npoint, a.merge_mlp[2], nullptr, m->motion,
                                     (int)align_up(3 + a.merge_mlp[2], 4), st));
        }
    }
    return run_head(m, B, y_out, st);
}

extern "C" int dclr_model_last_launches(const dclr_model *m) { return m ? m->last_launches : 0; }
extern "C" int dclr_model_generation(const dclr_model *m) { return m ? m->ws_generation : -1; }

extern "C" int dclr_model_set_timing(dclr_model *m, int on) {
    DCLR_CHECK_ARG(m, "null model");
    if (on && !m->ev) {
        m->ev = new (std::nothrow) cudaEvent_t[(size_t)EV_RING * EV_PER_CALL];
        DCLR_CHECK_ARG(m->ev, "out of host memory");
        for (int i = 0; i < EV_RING * EV_PER_CALL; ++i) DCLR_CUDA(cudaEventCreate(&m->ev[i]));
    }
    m->timing = on != 0;
    m->ev_calls = 0;
    return DCLR_OK;
}

extern "C" int dclr_model_stage_count(void) { return ST_COUNT; }
extern "C" const char *dclr_model_stage_name(int s) { return (s >= 0 && s < ST_COUNT) ? kStageNames[s] : ""; }

extern "C" int dclr_model_collect_timing(dclr_model *m, float *stage_ms_sum, int *calls_out) {
    DCLR_CHECK_ARG(m && stage_ms_sum && calls_out, "dclr_model_collect_timing: null pointer");
    DCLR_CHECK_ARG(m->ev, "dclr_model_collect_timing: timing was never enabled");
    for (int s = 0; s < ST_COUNT; ++s) stage_ms_sum[s] = 0.f;
    const int n = m->ev_calls < EV_RING ? m->ev_calls : EV_RING;
    for (int c = 0; c < n; ++c)
        for (int s = 0; s < ST_COUNT; ++s) {
            if (!(m->ev_mask[c] >> s & 1u)) continue;
            cudaEvent_t e0 = m->ev[(size_t)c * EV_PER_CALL + 2 * s], e1 = m->ev[(size_t)c * EV_PER_CALL + 2 * s + 1];
            DCLR_CUDA(cudaEventSynchronize(e1));
            float ms = 0.f;
            DCLR_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            stage_ms_sum[s] += ms;
        }
    *calls_out = n;
    m->ev_calls = 0;
    return DCLR_OK;
}

extern "C" int dclr_model_debug_copy(const dclr_model *m, int which, void *dst, size_t nbytes, void *stream) {
    DCLR_CHECK_ARG(m && m->ws && dst, "dclr_model_debug_copy: no workspace yet or null dst");
    const void *src = nullptr;
    switch (which) {
        case 0: src = m->fps_idx; break;
        case 1: src = m->bq_cnt[0]; break;
        case 2: src = m->bq_cnt[1]; break;
        case 3: src = m->knn_col; break;
        case 4: src = m->bq_idx[0]; break;
        case 5: src = m->bq_idx[1]; break;
        default: return fail(DCLR_E_INVALID, "dclr_model_debug_copy: unknown tap %d", which);
    }
    DCLR_CUDA(cudaMemcpyAsync(dst, src, nbytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return DCLR_OK;
}
