// Device-side pieces of the hash-grid encode kernels shared by lb_encode.cu and the fused backward in
// lb_mlp_umma.cu: sample-point construction, per-level cell / index math for one lane's x-side, and the
// per-point backward body (table scatter-add with warp-aggregated atomics, optional dL/dx).  See lb_encode.cu
// for the mapping (two lanes per sample point).
#pragma once
#include "lb_common.cuh"

namespace lb_enc {

constexpr unsigned kFull = 0xffffffffu;

struct Point {
    float x, y, z;
};

template <bool kFromRays>
__device__ __forceinline__ Point load_point(const float* __restrict__ rays, const float* __restrict__ zv,
                                            const float* __restrict__ x01, size_t p, int n_samples, float* z_out) {
    Point q;
    if (kFromRays) {
        const uint32_t r = (uint32_t)p / (uint32_t)n_samples;     // P < 2^32 (checked by the launchers)
        const float* ray = rays + (size_t)r * LB_RAY_STRIDE;
        const float z = __ldcs(zv + p);                            // use-once stream: evict-first, the tables keep the L2
        // same roundings as the reference's torch ops (o + d*z, then (xyz+1)/2): at the finest level one
        // ulp of x01 is 3 % of a cell, so the sample position must match bit for bit
        q.x = __fmul_rn(__fadd_rn(__fadd_rn(__ldg(ray + 0), __fmul_rn(__ldg(ray + 3), z)), 1.0f), 0.5f);
        q.y = __fmul_rn(__fadd_rn(__fadd_rn(__ldg(ray + 1), __fmul_rn(__ldg(ray + 4), z)), 1.0f), 0.5f);
        q.z = __fmul_rn(__fadd_rn(__fadd_rn(__ldg(ray + 2), __fmul_rn(__ldg(ray + 5), z)), 1.0f), 0.5f);
        if (z_out) *z_out = z;
    } else {
        q.x = __ldcs(x01 + 3 * p + 0);
        q.y = __ldcs(x01 + 3 * p + 1);
        q.z = __ldcs(x01 + 3 * p + 2);
    }
    return q;
}

// slow path (out-of-cube points on dense levels, non-power-of-two hashed levels): kept out of line so that the
// sixteen unrolled level bodies stay small enough for the instruction cache
static __device__ __noinline__ uint4 wrap_indices(uint4 i, uint32_t size) {
    return make_uint4(i.x % size, i.y % size, i.z % size, i.w % size);
}

// One lane's view of a level: cell, fractional position, its x-side weight and the entry indices of its four
// (y,z) corners (k = by + 2*bz), level offset included.
struct Side {
    uint32_t idx[4];
    float wx, wy[2], wz[2];
    uint32_t cx, cy, cz;         // cell (for run detection)
};

template <bool kDense>
__device__ __forceinline__ Side locate_side_t(const lb_level& L, const Point& q, int half) {
    Side s;
    const float px = fmaf(L.scale, q.x, 0.5f), py = fmaf(L.scale, q.y, 0.5f), pz = fmaf(L.scale, q.z, 0.5f);
    const float flx = floorf(px), fly = floorf(py), flz = floorf(pz);
    s.cx = (uint32_t)(int)flx; s.cy = (uint32_t)(int)fly; s.cz = (uint32_t)(int)flz;
    const float fx = px - flx, fy = py - fly, fz = pz - flz;
    s.wx = half ? fx : 1.0f - fx;
    s.wy[0] = 1.0f - fy; s.wy[1] = fy;
    s.wz[0] = 1.0f - fz; s.wz[1] = fz;
    const uint32_t x = s.cx + (uint32_t)half;
    if (kDense) {
        // x + y*res + z*res^2; tcnn wraps with "% size", which in-cube points never need
        const uint32_t a0 = x + s.cy * L.res, a1 = a0 + L.res;
        const uint32_t z0 = s.cz * L.res2, z1 = z0 + L.res2;
        s.idx[0] = a0 + z0; s.idx[1] = a1 + z0; s.idx[2] = a0 + z1; s.idx[3] = a1 + z1;
        if (s.idx[3] >= L.size || s.idx[0] >= L.size) {
            const uint4 w = wrap_indices(make_uint4(s.idx[0], s.idx[1], s.idx[2], s.idx[3]), L.size);
            s.idx[0] = w.x; s.idx[1] = w.y; s.idx[2] = w.z; s.idx[3] = w.w;
        }
    } else {
        // coherent prime hash; a hashed level always has size = 2^log2_hashmap_size (lb_plan_create), so the
        // "% size" of tcnn is a mask
        const uint32_t y0 = s.cy * 2654435761u, y1 = y0 + 2654435761u;
        const uint32_t z0 = s.cz * 805459861u, z1 = z0 + 805459861u;
        s.idx[0] = (x ^ y0 ^ z0) & L.mask; s.idx[1] = (x ^ y1 ^ z0) & L.mask;
        s.idx[2] = (x ^ y0 ^ z1) & L.mask; s.idx[3] = (x ^ y1 ^ z1) & L.mask;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) s.idx[k] += L.offset;
    return s;
}

// kNDense >= 0: levels [0, kNDense) are dense and the rest hashed, resolved at compile time once the level loop
// is unrolled (one index formula per level body instead of two); kNDense < 0: decided per level at run time
template <int kNDense>
__device__ __forceinline__ Side locate_side(const lb_level& L, const Point& q, int half, int l) {
    if (kNDense >= 0) return (l < kNDense) ? locate_side_t<true>(L, q, half) : locate_side_t<false>(L, q, half);
    return L.dense ? locate_side_t<true>(L, q, half) : locate_side_t<false>(L, q, half);
}

// One lane's forward: gathers the 4 (y,z) corners of its x-side on every level in [l_begin, l_end), then the pair
// exchange: the even lane ends up with features 0-15 of the point, the odd lane with features 16-31 (zeros for
// levels outside the range).  Called convergently by both lanes of a pair.
template <int kNDense>
__device__ __forceinline__ void encode_fwd_point(const lb_level_table& T, const Point& q, int half,
                                                 const float2* __restrict__ table, int l_begin, int l_end, float* out) {
    float part[LB_MLP_IN];
#pragma unroll
    for (int l = 0; l < 16; ++l) {
        float a = 0.0f, b = 0.0f;
        if (l >= l_begin && l < l_end) {                       // level-group pass (whole range when the tables fit L2)
            const lb_level& L = T.lv[l];
            const Side s = locate_side<kNDense>(L, q, half, l);
            float2 v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = __ldg(table + s.idx[k]);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float w = s.wy[k & 1] * s.wz[k >> 1];
                a = fmaf(w, v[k].x, a);
                b = fmaf(w, v[k].y, b);
            }
            a *= s.wx;
            b *= s.wx;
        }
        part[2 * l] = a;
        part[2 * l + 1] = b;
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float send = half ? part[i] : part[16 + i];
        const float recv = __shfl_xor_sync(kFull, send, 1);
        out[i] = (half ? part[16 + i] : part[i]) + recv;
    }
}

// encode_fwd_point plus the spatial Jacobian of the encoding, d e[2l+f] / d x01[d] (fused forward with forward-mode
// input derivative, lb_mlp_umma.cu).  Per level and x-side a lane holds the four (y,z) corners v[k], so
//   d/dx = scale * (A(x+1) - A(x)),               A = sum_k wy wz v[k]   (the e value before the x weight)
//   d/dy = scale * wx * (wz0 (v1 - v0) + wz1 (v3 - v2)),  d/dz likewise, summed over the two x-sides.
// Pair exchange as for e: the even lane ends up with the three derivatives of feature 0, the odd lane with those of
// feature 1, and hands them to `store(g, jl)` four levels at a time (jl[(l & 3) * 3 + d], g = l / 4: 12 floats, so that
// the Jacobian never occupies more than 12 registers).  The destination -- a row in shared memory or columns of the
// lane in tensor memory -- is the caller's; its first call (g = 0, after the arithmetic of levels 0-3) is where the
// caller waits for the destination to become free, so a quarter of the point's gathers overlap that wait.
template <int kNDense, typename StoreGroup>
__device__ __forceinline__ void encode_fwd_jvp_point(const lb_level_table& T, const Point& q, int half,
                                                     const float2* __restrict__ table, float* out, StoreGroup store) {
    float part[LB_MLP_IN];
    float jl[12];
#pragma unroll
    for (int l = 0; l < 16; ++l) {
        const lb_level& L = T.lv[l];
        const Side s = locate_side<kNDense>(L, q, half, l);
        float2 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = __ldg(table + s.idx[k]);
        float a = 0.0f, b = 0.0f;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float w = s.wy[k & 1] * s.wz[k >> 1];
            a = fmaf(w, v[k].x, a);
            b = fmaf(w, v[k].y, b);
        }
        const float ya = s.wx * fmaf(s.wz[1], v[3].x - v[2].x, s.wz[0] * (v[1].x - v[0].x));
        const float yb = s.wx * fmaf(s.wz[1], v[3].y - v[2].y, s.wz[0] * (v[1].y - v[0].y));
        const float za = s.wx * fmaf(s.wy[1], v[3].x - v[1].x, s.wy[0] * (v[2].x - v[0].x));
        const float zb = s.wx * fmaf(s.wy[1], v[3].y - v[1].y, s.wy[0] * (v[2].y - v[0].y));
        part[2 * l] = a * s.wx;
        part[2 * l + 1] = b * s.wx;
        // the even lane (x side) owns feature 0 and sends its feature-1 partials; the odd lane (x + 1 side) the reverse
        const float r0 = __shfl_xor_sync(kFull, half ? a : b, 1);
        const float r1 = __shfl_xor_sync(kFull, half ? ya : yb, 1);
        const float r2 = __shfl_xor_sync(kFull, half ? za : zb, 1);
        const int g = (l & 3) * 3;
        jl[g + 0] = L.scale * (half ? b - r0 : r0 - a);
        jl[g + 1] = L.scale * ((half ? yb : ya) + r1);
        jl[g + 2] = L.scale * ((half ? zb : za) + r2);
        if ((l & 3) == 3) store(l >> 2, jl);
    }
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float send = half ? part[i] : part[16 + i];
        const float recv = __shfl_xor_sync(kFull, send, 1);
        out[i] = (half ? part[16 + i] : part[i]) + recv;
    }
}

// ------------------------------------------------------------------------------------------
// backward: table scatter-add (warp-aggregated) and optional dL/dx01
// ------------------------------------------------------------------------------------------
// Run merging of the backward scatter: a level below kAggLevels merges the warp's 16 points into runs of equal cells
// when they fall into <= kAggMaxRuns distinct cells; only those levels carry the merging code (instruction-cache
// budget).  Tuned per kernel on uniform and surface-concentrated depths (profiles/README.md, "run-merging sweep"):
// the stand-alone scatter kernel is fastest with (8, 15), the fused MLP + scatter kernel -- timed inside the bench
// step, whose d_sigma has the real sparsity -- with (7, 8).
constexpr int kAggLevelsScatter = 8, kAggMaxRunsScatter = 15;
#ifndef LB_AGG_LEVELS_FUSED
#define LB_AGG_LEVELS_FUSED 7
#define LB_AGG_MAX_RUNS_FUSED 8
#endif
constexpr int kAggLevelsFused = LB_AGG_LEVELS_FUSED, kAggMaxRunsFused = LB_AGG_MAX_RUNS_FUSED;

// segmented suffix sum over the warp's 16 points (2 lanes each): a run head collects its whole run.  Out of line (one
// copy instead of one per level), called convergently by all 32 lanes, and on all eight values of a level at once:
// the four shuffle rounds are a dependent chain (~100 cycles), eight independent chains overlap instead of queueing.
// Measured: the batched form wins when the kernel also computes dL/dx (4.08 -> 3.70 ms), the one-value form wins in
// the table-only kernel (2.39 vs 2.83 ms, instruction-cache footprint); kInputGrad selects.
static __device__ __noinline__ float seg_suffix_sum(float v, int dist) {
#pragma unroll
    for (int off = 1; off < 16; off <<= 1) {
        const float t = __shfl_down_sync(kFull, v, 2 * off);
        if (off < dist) v += t;
    }
    return v;
}
struct Float8 {
    float v[8];
};
static __device__ __noinline__ Float8 seg_suffix_sum8(Float8 x, int dist) {
#pragma unroll
    for (int off = 1; off < 16; off <<= 1) {
        float t[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) t[i] = __shfl_down_sync(kFull, x.v[i], 2 * off);
        const float m = off < dist ? 1.0f : 0.0f;      // branch-free: lanes of one warp have different run lengths
#pragma unroll
        for (int i = 0; i < 8; ++i) x.v[i] = fmaf(m, t[i], x.v[i]);
    }
    return x;
}

// One point's backward given its upstream gradient row de[32] (zero outside [l_begin, l_end)).  Called convergently
// by all 32 lanes of a warp (16 points x 2 lanes); lane parity = x-side.
template <bool kFromRays, bool kInputGrad, int kNDense, int kAggLevels = kAggLevelsScatter, int kAggMaxRuns = kAggMaxRunsScatter>
__device__ __forceinline__ void encode_bwd_point(const lb_level_table& T, const Point& q, float zval, const float* de,
                                                 bool valid, size_t p, int n_samples, const float2* __restrict__ table,
                                                 float* __restrict__ grad_table, float* __restrict__ d_rays,
                                                 float* __restrict__ d_x01, int l_begin, int l_end) {
    const int lane = threadIdx.x & 31;
    const int half = lane & 1;
    const int pair = lane >> 1;
    // A sample whose upstream gradient row is exactly zero (ReLU-inactive density: dL/dsigma = 0, roughly
    // every second sample while the field is young) scatters nothing: the L2 atomics are what bounds this
    // kernel, so such points issue none.
    uint32_t nz_bits = 0u;
#pragma unroll
    for (int i = 0; i < LB_MLP_IN; ++i) nz_bits |= __float_as_uint(de[i]);
    const bool nonzero = (nz_bits << 1) != 0u;
    if (__all_sync(kFull, !nonzero)) {                             // warp-uniform: nothing to scatter for these 16 points
        if (kInputGrad && !kFromRays && valid && !half && l_begin == 0) { d_x01[3 * p + 0] = 0.0f; d_x01[3 * p + 1] = 0.0f; d_x01[3 * p + 2] = 0.0f; }
        return;
    }
    float gx = 0.0f, gy = 0.0f, gz = 0.0f;
    bool all_distinct = false;     // once a level has 16 distinct cells, finer levels do too
#pragma unroll
    for (int l = 0; l < 16; ++l) {
        if (l < l_begin || l >= l_end) continue;
        const lb_level& L = T.lv[l];
        const Side s = locate_side<kNDense>(L, q, half, l);
        const float d0 = de[2 * l], d1 = de[2 * l + 1];
        const uint32_t* idx = s.idx;
        if (kInputGrad) {
            float val[4], oth[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 v = __ldg(table + idx[k]);
                val[k] = fmaf(v.x, d0, v.y * d1);
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) oth[k] = __shfl_xor_sync(kFull, val[k], 1);
            float tx = 0.0f;
#pragma unroll
            for (int k = 0; k < 4; ++k) tx = fmaf(s.wy[k & 1] * s.wz[k >> 1], oth[k] - val[k], tx);
            if (half) tx = 0.0f;                       // the even lane holds (x+1) - (x)
            const float ty = s.wx * (s.wz[0] * (val[1] - val[0]) + s.wz[1] * (val[3] - val[2]));
            const float tz = s.wx * (s.wy[0] * (val[2] - val[0]) + s.wy[1] * (val[3] - val[1]));
            gx = fmaf(L.scale, tx, gx);
            gy = fmaf(L.scale, ty, gy);
            gz = fmaf(L.scale, tz, gz);
        }
        if (grad_table) {
            float a[4], b[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float w = s.wx * s.wy[k & 1] * s.wz[k >> 1];
                a[k] = w * d0;
                b[k] = w * d1;
            }
            bool issue = valid && nonzero;
            if (l < kAggLevels && !all_distinct) {
                // run detection over the warp's 16 points (previous point = lane - 2)
                const uint32_t pcx = __shfl_up_sync(kFull, s.cx, 2), pcy = __shfl_up_sync(kFull, s.cy, 2),
                               pcz = __shfl_up_sync(kFull, s.cz, 2);
                const int pvalid = __shfl_up_sync(kFull, (int)valid, 2);
                const bool same = (lane >= 2) && valid && pvalid && pcx == s.cx && pcy == s.cy && pcz == s.cz;
                const unsigned heads = __ballot_sync(kFull, !same) & 0x55555555u;
                const int runs = __popc(heads);
                if (runs == 16) {
                    all_distinct = true;
                } else if (runs <= kAggMaxRuns) {
                    // segmented suffix sum: the head point of each run collects the whole run
                    const unsigned rest = (pair < 15) ? (heads >> (2 * pair + 2)) : 0u;
                    const int dist = rest ? 1 + ((__ffs(rest) - 1) >> 1) : 16 - pair;   // points until the next head
                    if (kInputGrad) {
                        Float8 x;
#pragma unroll
                        for (int k = 0; k < 4; ++k) { x.v[k] = a[k]; x.v[4 + k] = b[k]; }
                        x = seg_suffix_sum8(x, dist);
#pragma unroll
                        for (int k = 0; k < 4; ++k) { a[k] = x.v[k]; b[k] = x.v[4 + k]; }
                    } else {
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            a[k] = seg_suffix_sum(a[k], dist);
                            b[k] = seg_suffix_sum(b[k], dist);
                        }
                    }
                    issue = valid && !same;      // a run head carries the run's sum (zero rows add zeros)
                }
            }
            if (issue) {
#pragma unroll
                for (int k = 0; k < 4; ++k) red_add_f32x2(grad_table + 2 * (size_t)idx[k], a[k], b[k]);
            }
        }
    }
    if (kInputGrad) {
        gx += __shfl_xor_sync(kFull, gx, 1);
        gy += __shfl_xor_sync(kFull, gy, 1);
        gz += __shfl_xor_sync(kFull, gz, 1);
        if (kFromRays) {
            // d_o += G/2, d_d += z*G/2; one set of atomics per warp when it sits inside one ray
            const bool mine = valid && !half;
            float v[6] = {0.5f * gx, 0.5f * gy, 0.5f * gz, 0.5f * zval * gx, 0.5f * zval * gy, 0.5f * zval * gz};
            if (!mine) {
#pragma unroll
                for (int i = 0; i < 6; ++i) v[i] = 0.0f;
            }
            const unsigned long long ray = valid ? (unsigned long long)((uint32_t)p / (uint32_t)n_samples) : ~0ull;
            const unsigned long long r0 = __shfl_sync(kFull, ray, 0);
            const bool uniform = __all_sync(kFull, !valid || ray == r0) && (r0 != ~0ull);
            if (uniform) {
#pragma unroll
                for (int i = 0; i < 6; ++i) v[i] = warp_sum(v[i]);
                if (lane == 0) {
#pragma unroll
                    for (int i = 0; i < 6; ++i) atomicAdd(d_rays + r0 * LB_RAY_STRIDE + i, v[i]);
                }
            } else if (mine) {
#pragma unroll
                for (int i = 0; i < 6; ++i) atomicAdd(d_rays + ray * LB_RAY_STRIDE + i, v[i]);
            }
        } else if (valid && !half) {
            if (l_begin == 0) {
                d_x01[3 * p + 0] = gx; d_x01[3 * p + 1] = gy; d_x01[3 * p + 2] = gz;
            } else {                                   // later level-group passes accumulate
                d_x01[3 * p + 0] += gx; d_x01[3 * p + 1] += gy; d_x01[3 * p + 2] += gz;
            }
        }
    }
}

}  // namespace lb_enc
