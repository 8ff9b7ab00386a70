// Pre-processing in front of the forward, batched on the device over a RAGGED batch of raw scans:
//   filter_clouds_kernel   SystematicErasing (cloud[start::nth], deepclr/data/transforms/transforms.py:244-268)
//                          -> RangeSelection (max(|x|,|y|) in [min,max], :90-113) -> TruncateDimension (:271-283),
//                          the order of deepclr/data/transforms/build.py:23-24,37-38; order-preserving compaction
//   subsample_clouds_kernel size-equalising random subsample into the dense [num_clouds, N, C] batch the forward reads
//                          (ModelInferenceHelper.stack, deepclr/models/base.py:130-135: cloud[randperm(n)[:N]];
//                          RandomErasing max_points, transforms.py:128-131)
// The reference does all of this per cloud in numpy on the host and then copies the result; here the raw scans are
// uploaded once and never come back.
//
// Both kernels are HBM-bound byte movers.  The filter is ONE pass: a tile of 2048 points is loaded (one float4 per
// point), its keep flags are ranked with ballots + a 64-segment scan, and the tile's place in the output comes from a
// decoupled look-back over the tiles of its cloud (one 64-bit status word per tile: own count, later inclusive prefix;
// warp 0 inspects 32 predecessors per round trip).  Tiles are handed out by an atomic ticket, so every tile a CTA looks
// back at is held by a CTA that is already running.  (First version: a chain {next tile, total} per cloud -- correct,
// but 65 serial L2 round trips per 120 k-point scan with ~18 scans in flight: 0.47 ms per 128 scans, 9.6 % of HBM.)
// Algorithmic bytes: 4 C_in n read + 4 C_out kept written.
#include "common.cuh"
#include "kernels.cuh"

namespace dclr {

constexpr int FLT_THREADS = 256;
#ifndef DCLR_FLT_ROWS
#define DCLR_FLT_ROWS 8
#endif
constexpr int FLT_ROWS = DCLR_FLT_ROWS;            // rows of 256 points per tile, read 8 at a time
static_assert(FLT_ROWS % 8 == 0 && FLT_ROWS <= 32, "tile height: a multiple of 8 rows, at most 32");
constexpr int FLT_TILE = FLT_THREADS * FLT_ROWS;   // 2048 points = 32 KB of float4
constexpr int FLT_SEGS = FLT_ROWS * (FLT_THREADS / 32);
constexpr int FLT_PER_LANE = FLT_SEGS / 32;

struct FilterArgs {
    int nth, start;          // keep i iff i >= start && (i - start) % nth == 0
    float rmin, rmax;        // then keep iff rmin <= max(|x|,|y|) <= rmax (fp32 compare, NaN coordinates drop)
    int use_range;           // the reference returns the cloud untouched for (0, inf)
};

__device__ __forceinline__ bool keep_point(int i, float x, float y, const FilterArgs &f) {
    bool k = true;
    if (f.nth > 1) k = (i >= f.start) && ((i - f.start) % f.nth == 0);
    else k = i >= f.start;
    if (f.use_range) {
        const float m = fmaxf(fabsf(x), fabsf(y));
        k = k && (x == x) && (y == y) && (m >= f.rmin) && (m <= f.rmax);   // np.max propagates NaN -> both tests false
    }
    return k;
}

__device__ __forceinline__ unsigned long long ld_chain(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_chain(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Tile status word: flag in the high half (0 = nothing yet, 1 = the tile's own count, 2 = inclusive prefix of the
// cloud up to and including this tile), the number of points in the low half.
constexpr unsigned long long ST_AGG = 1ull << 32, ST_PREFIX = 2ull << 32;

// grid (max_tiles, B), one tile per CTA.  Tiles are handed out by a per-cloud ticket, not by blockIdx: every tile a CTA
// looks back at was taken earlier, i.e. by a CTA that is running or done, so the waits in the look-back cannot starve
// whatever order the CTAs are scheduled in.  VEC4: C_in == C_out == 4, rows are float4.
// (Tried: persistent CTAs drawing (cloud, tile) pairs from ONE global counter with the next ticket PREFETCHED at the start
// of the current tile -- 0.79 ms instead of 0.099 per 128 scans, growing with the number of resident CTAs: a prefetched
// ticket is a tile nobody works on yet, so a look-back can wait for a CTA that first has to finish its own tile, whose
// look-back waits for another such CTA ...: dependency chains as long as the grid.  A ticket must be drawn at the
// moment its tile is started.  Taller tiles (16 / 24 / 32 rows) were slower as well: 0.115 / 0.117 / 0.178 ms.)
#ifndef DCLR_FLT_MINB
#define DCLR_FLT_MINB 8
#endif
#ifndef DCLR_FLT_PREFETCH
#define DCLR_FLT_PREFETCH 1
#endif
template <bool VEC4>
__global__ void __launch_bounds__(FLT_THREADS, DCLR_FLT_MINB)
filter_clouds_kernel(const float *__restrict__ pts, const int32_t *__restrict__ offsets, int B, int C_in, int C_out,
                     FilterArgs f, float *__restrict__ out, int32_t *__restrict__ counts,
                     unsigned *__restrict__ ticket, unsigned long long *__restrict__ status, int max_tiles) {
    __shared__ int s_seg[FLT_SEGS];
    __shared__ unsigned s_ticket;
    __shared__ int s_base;
    const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
    const int b = blockIdx.y;
    const int row0 = offsets[b];
    const int n = offsets[b + 1] - row0;
    const int ntiles = (n + FLT_TILE - 1) / FLT_TILE;
    if (threadIdx.x == 0) s_ticket = atomicAdd(&ticket[b], 1u);
#if DCLR_FLT_PREFETCH
    if ((int)blockIdx.x < ntiles && ntiles <= max_tiles) {
        // While the ticket is on its way: request tile blockIdx.x from DRAM (fire-and-forget L2 prefetches, one per
        // 128-byte line, no registers held).  CTAs start in blockIdx order, so the ticket is usually this very tile; if
        // not, the tile belongs to another CTA of this launch and the prefetch serves that one.
        const char *tile = reinterpret_cast<const char *>(pts + ((size_t)row0 + (size_t)blockIdx.x * FLT_TILE) * C_in);
        const size_t bytes = (size_t)min(FLT_TILE, n - (int)blockIdx.x * FLT_TILE) * C_in * sizeof(float);
        for (size_t o = (size_t)threadIdx.x * 128; o < bytes; o += (size_t)FLT_THREADS * 128)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(tile + o));
    }
#endif
    __syncthreads();
    const unsigned t = s_ticket;
    {
        if (ntiles > max_tiles || (int)t >= ntiles) {
            // cloud longer than the caller's max_points: reported, not run; empty cloud: count 0
            if (t == 0 && threadIdx.x == 0 && (ntiles > max_tiles || ntiles == 0)) counts[b] = ntiles ? -1 : 0;
            return;
        }
        unsigned long long *st = status + (size_t)b * max_tiles;
        const int i0 = (int)t * FLT_TILE + (int)threadIdx.x;

        // phase 1 reads only x, y (the sectors arrive whole and stay in L1 / L2); the surviving rows are read again in
        // phase 3.  Holding the float4 rows instead costs 64 registers = 4 CTAs per SM, and the CTAs that sit in their
        // look-back then leave the memory system idle (measured: 0.126 vs 0.099 ms per 128 scans).
        unsigned keep_bits = 0, rank_bits[FLT_ROWS / 4];   // per row: the keep flag, the rank inside its warp (8 bits)
#pragma unroll
        for (int q = 0; q < FLT_ROWS / 4; ++q) rank_bits[q] = 0;
#pragma unroll
        for (int r0 = 0; r0 < FLT_ROWS; r0 += 8) {
            float2 xy[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int i = i0 + (r0 + r) * FLT_THREADS;
                xy[r] = make_float2(0.f, 0.f);
                if (i < n) {
                    if (VEC4) {
                        xy[r] = __ldg(reinterpret_cast<const float2 *>(pts + (size_t)(row0 + i) * 4));
                    } else {
                        const float *p = pts + (size_t)(row0 + i) * C_in;
                        xy[r] = make_float2(__ldg(p), __ldg(p + 1));
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const int i = i0 + (r0 + r) * FLT_THREADS;
                const bool k = (i < n) && keep_point(i, xy[r].x, xy[r].y, f);
                const unsigned bal = __ballot_sync(0xffffffffu, k);
                if (lane == 0) s_seg[(r0 + r) * (FLT_THREADS / 32) + warp] = __popc(bal);
                const unsigned rk = __popc(bal & ((1u << lane) - 1u));
                keep_bits |= (k ? 1u : 0u) << (r0 + r);
                rank_bits[(r0 + r) >> 2] |= rk << (((r0 + r) & 3) * 8);
            }
        }
        __syncthreads();
        if (warp == 0) {
            // exclusive scan of the segment counts (row-major: all warps of row 0, then row 1 ...), FLT_PER_LANE per lane
            int c[FLT_PER_LANE], sum = 0;
#pragma unroll
            for (int q = 0; q < FLT_PER_LANE; ++q) { c[q] = s_seg[FLT_PER_LANE * lane + q]; sum += c[q]; }
            int incl = sum;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int o = __shfl_up_sync(0xffffffffu, incl, d);
                if ((int)lane >= d) incl += o;
            }
            int run = incl - sum;
#pragma unroll
            for (int q = 0; q < FLT_PER_LANE; ++q) { s_seg[FLT_PER_LANE * lane + q] = run; run += c[q]; }
            const unsigned total = (unsigned)__shfl_sync(0xffffffffu, incl, 31);
            // decoupled look-back over the earlier tiles of this cloud: publish the own count at once, then add up the
            // counts of the predecessors back to the nearest one that already knows its prefix, 32 tiles per round trip
            if (lane == 0) st_chain(st + t, (t == 0 ? ST_PREFIX : ST_AGG) | total);
            unsigned prefix = 0;
            if (t > 0) {
                int look = (int)t - 1;
                for (;;) {
                    // 64 predecessors per round trip: lane l inspects tiles look - 2l and look - 2l - 1 ("tile -1" has
                    // prefix 0); a KITTI scan has <= 65 tiles, so one poll usually settles it
                    const int i0 = look - 2 * (int)lane, i1 = i0 - 1;
                    const unsigned long long w0 = i0 >= 0 ? ld_chain(st + i0) : ST_PREFIX;
                    const unsigned long long w1 = i1 >= 0 ? ld_chain(st + i1) : ST_PREFIX;
                    const unsigned f0 = (unsigned)(w0 >> 32), f1 = (unsigned)(w1 >> 32);
                    const unsigned mypos = f0 == 2u ? 2u * lane : (f1 == 2u ? 2u * lane + 1u : 64u);
                    const unsigned fp = __reduce_min_sync(0xffffffffu, mypos);     // nearest tile with a prefix
                    const bool bad = (2u * lane <= fp && f0 == 0u) || (2u * lane + 1u <= fp && f1 == 0u);
                    if (__any_sync(0xffffffffu, bad)) {                            // a needed tile has not published yet
                        __nanosleep(20);
                        continue;
                    }
                    const unsigned val = (2u * lane <= fp ? (unsigned)w0 : 0u) + (2u * lane + 1u <= fp ? (unsigned)w1 : 0u);
                    prefix += __reduce_add_sync(0xffffffffu, val);
                    if (fp < 64u) break;
                    look -= 64;
                }
                if (lane == 0) st_chain(st + t, ST_PREFIX | (unsigned long long)(prefix + total));
            }
            if (lane == 0) {
                s_base = (int)prefix;
                if ((int)t == ntiles - 1) counts[b] = (int)(prefix + total);
            }
        }
        __syncthreads();
        const int base = s_base;
#pragma unroll
        for (int r = 0; r < FLT_ROWS; ++r) {
            if (!((keep_bits >> r) & 1u)) continue;
            const int rk = (rank_bits[r >> 2] >> ((r & 3) * 8)) & 0xff;
            const size_t dst = (size_t)row0 + base + s_seg[r * (FLT_THREADS / 32) + warp] + rk;
            if (VEC4) {
                __stcs(reinterpret_cast<float4 *>(out) + dst,
                       __ldg(reinterpret_cast<const float4 *>(pts) + (size_t)(row0 + i0 + r * FLT_THREADS)));
            } else {
                const float *p = pts + (size_t)(row0 + i0 + r * FLT_THREADS) * C_in;
                for (int c = 0; c < C_out; ++c) out[dst * C_out + c] = __ldg(p + c);
            }
        }
    }
}

// ---- keyed permutation of [0, n): 4-round Feistel network on ceil(log2 n) bits, cycle walking back into the domain
// (first version: 6 rounds on an even bit count, up to 4 walks + an integer division per row -- the kernel was bound by
// the integer pipe, 0.072 ms per 128 x 16384 rows).  A bijection by construction, so the first N values are N distinct rows.
__device__ __forceinline__ unsigned mix32(unsigned x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ unsigned feistel_perm(unsigned j, unsigned n, unsigned key) {
    if (n <= 1) return 0;
    int bits = 32 - __clz(n - 1);
    if (bits < 2) bits = 2;
    // L = the high `ha` bits, R = the low `hb` bits, ha >= hb (unbalanced for an odd bit count: the domain stays
    // below 2n, < 2 walks on average); a round maps (L, R) -> (R, L ^ F(R)) and swaps the two widths, an even number
    // of rounds restores them
    const int hb = bits >> 1, ha = bits - hb;
    const unsigned ma = (1u << ha) - 1u, mb = (1u << hb) - 1u;
    unsigned x = j;
    do {
        unsigned L = x >> hb, R = x & mb;
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const unsigned F = mix32(R ^ (key + 0x9e3779b9u * (unsigned)(r + 1))) & ((r & 1) ? mb : ma);
            const unsigned nl = R;
            R = L ^ F;
            L = nl;
        }
        x = (L << hb) | R;
    } while (x >= n);
    return x;
}

// grid (ceil(N_out / (256 * SUB_ROWS)), B): SUB_ROWS output rows per thread, all gathers in flight before the stores
constexpr int SUB_ROWS = 4;
template <bool VEC4>
__global__ void __launch_bounds__(256)
subsample_clouds_kernel(const float *__restrict__ in, const int32_t *__restrict__ offsets,
                        const int32_t *__restrict__ counts, int C_in, int C_out, int N_out, unsigned seed_lo,
                        unsigned seed_hi, int ordered_if_equal, float *__restrict__ x_out, int32_t *__restrict__ perm_out,
                        int32_t *__restrict__ short_out) {
    const int b = blockIdx.y;
    const int j0 = blockIdx.x * (256 * SUB_ROWS) + threadIdx.x;
    const int row0 = offsets[b];
    const int n = counts ? counts[b] : offsets[b + 1] - row0;
    if (n < N_out && j0 == 0 && short_out) atomicAdd(short_out, 1);
    const unsigned key = mix32(seed_lo ^ mix32(seed_hi + (unsigned)b));
    const bool ordered = (n == N_out) && ordered_if_equal;   // the reference leaves equal-sized clouds in order
    int src[SUB_ROWS];
    float4 v[SUB_ROWS];
#pragma unroll
    for (int r = 0; r < SUB_ROWS; ++r) {
        const int j = j0 + r * 256;
        src[r] = -1;
        if (j < N_out && n > 0) src[r] = ordered ? j : (int)feistel_perm((unsigned)(j < n ? j : j % n), (unsigned)n, key);
        v[r] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (VEC4 && src[r] >= 0) v[r] = __ldg(reinterpret_cast<const float4 *>(in) + (size_t)(row0 + src[r]));
    }
#pragma unroll
    for (int r = 0; r < SUB_ROWS; ++r) {
        const int j = j0 + r * 256;
        if (j >= N_out) continue;
        const size_t dst = (size_t)b * N_out + j;
        if (perm_out) perm_out[dst] = src[r];
        if (VEC4) {
            __stcs(reinterpret_cast<float4 *>(x_out) + dst, v[r]);
        } else {
            for (int c = 0; c < C_out; ++c)
                x_out[dst * C_out + c] = src[r] >= 0 ? __ldg(in + (size_t)(row0 + src[r]) * C_in + c) : 0.f;
        }
    }
}

}  // namespace dclr

using namespace dclr;

extern "C" int dclr_filter_clouds(const float *pts, const int32_t *offsets, int B, int max_points, int C_in, int C_out,
                                  int nth, int start, float min_range, float max_range, float *out,
                                  int32_t *counts_out, void *stream) {
    DCLR_CHECK_ARG(B >= 0 && max_points >= 0, "dclr_filter_clouds: B and max_points must be >= 0");
    if (B == 0) return DCLR_OK;
    DCLR_CHECK_ARG(pts && offsets && out && counts_out, "dclr_filter_clouds: null pointer");
    DCLR_CHECK_ARG(C_in >= 2 && C_out >= 1 && C_out <= C_in, "dclr_filter_clouds: need 2 <= C_in, 1 <= C_out <= C_in (got %d, %d)", C_in, C_out);
    DCLR_CHECK_ARG(nth >= 1 && start >= 0 && start < nth, "dclr_filter_clouds: need nth >= 1 and 0 <= start < nth (got %d, %d)", nth, start);
    DCLR_CHECK_ARG(B <= 65535, "dclr_filter_clouds: at most 65535 clouds per call");
    DCLR_CHECK_ARG(!(min_range != min_range) && !(max_range != max_range), "dclr_filter_clouds: NaN range");
    cudaStream_t st = (cudaStream_t)stream;
    FilterArgs f;
    f.nth = nth; f.start = start; f.rmin = min_range; f.rmax = max_range;
    f.use_range = !(min_range == 0.0f && isinf(max_range));   // np.isinf: either sign, as the reference tests it
    // per-tile status words (8 B) + per-cloud ticket (4 B), zeroed per call
    const int tiles = max_points > 0 ? div_up(max_points, FLT_TILE) : 1;
    void *ws = nullptr;
    const size_t chain_bytes = (size_t)B * tiles * 8, ws_bytes = chain_bytes + (size_t)B * 4;
    int rc = scratch_alloc(&ws, ws_bytes, st);
    if (rc != DCLR_OK) return rc;
    cudaError_t e = cudaMemsetAsync(ws, 0, ws_bytes, st);
    if (e != cudaSuccess) { scratch_free(ws, st); return fail(DCLR_E_CUDA, "dclr_filter_clouds: memset: %s", cudaGetErrorString(e)); }
    auto *chain = (unsigned long long *)ws;
    auto *ticket = (unsigned *)((char *)ws + chain_bytes);
    dim3 grid(tiles, B);
    const bool vec4 = C_in == 4 && C_out == 4 && ((uintptr_t)pts % 16 == 0) && ((uintptr_t)out % 16 == 0);
    if (vec4) filter_clouds_kernel<true><<<grid, FLT_THREADS, 0, st>>>(pts, offsets, B, C_in, C_out, f, out, counts_out, ticket, chain, tiles);
    else filter_clouds_kernel<false><<<grid, FLT_THREADS, 0, st>>>(pts, offsets, B, C_in, C_out, f, out, counts_out, ticket, chain, tiles);
    e = cudaGetLastError();
    scratch_free(ws, st);
    if (e != cudaSuccess) return fail(DCLR_E_CUDA, "launch of filter_clouds_kernel failed: %s", cudaGetErrorString(e));
    return DCLR_OK;
}

extern "C" int dclr_subsample_clouds(const float *pts, const int32_t *offsets, const int32_t *counts, int B, int C_in,
                                     int C_out, int N_out, uint64_t seed, int ordered_if_equal, float *x_out,
                                     int32_t *perm_out, int32_t *short_out, void *stream) {
    DCLR_CHECK_ARG(B >= 0 && N_out >= 0, "dclr_subsample_clouds: B and N_out must be >= 0");
    if (B == 0 || N_out == 0) return DCLR_OK;
    DCLR_CHECK_ARG(pts && offsets && x_out, "dclr_subsample_clouds: null pointer");
    DCLR_CHECK_ARG(C_in >= 1 && C_out >= 1 && C_out <= C_in, "dclr_subsample_clouds: need 1 <= C_out <= C_in (got %d, %d)", C_out, C_in);
    DCLR_CHECK_ARG(B <= 65535, "dclr_subsample_clouds: at most 65535 clouds per call");
    cudaStream_t st = (cudaStream_t)stream;
    if (short_out) DCLR_CUDA(cudaMemsetAsync(short_out, 0, sizeof(int32_t), st));
    dim3 grid(div_up(N_out, 256 * SUB_ROWS), B);
    const bool vec4 = C_in == 4 && C_out == 4 && ((uintptr_t)pts % 16 == 0) && ((uintptr_t)x_out % 16 == 0);
    const unsigned lo = (unsigned)seed, hi = (unsigned)(seed >> 32);
    if (vec4) subsample_clouds_kernel<true><<<grid, 256, 0, st>>>(pts, offsets, counts, C_in, C_out, N_out, lo, hi, ordered_if_equal, x_out, perm_out, short_out);
    else subsample_clouds_kernel<false><<<grid, 256, 0, st>>>(pts, offsets, counts, C_in, C_out, N_out, lo, hi, ordered_if_equal, x_out, perm_out, short_out);
    DCLR_LAUNCH_CHECK("subsample_clouds_kernel");
    return DCLR_OK;
}
