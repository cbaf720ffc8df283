// R4 (encoding half): multi-resolution hash-grid encode, forward and backward.
//
// Replaces tcnn.Encoding(HashGrid) / the encoding stage of tcnn.NetworkWithInputEncoding as used by
// DecoupledNeRF (reference src/models/nerf_tcnn.py:35-40,63-72), fused with the sample-point
// construction xyz = o + d*z (src/models/rendering_tcnn.py:390) and pos01 = (xyz+1)/2 (nerf_tcnn.py:63).
// Grid semantics follow the published tiny-cuda-nn algorithm (oracle/tinycudann/__init__.py).
//
// Mapping: TWO lanes per sample point.  Lane parity selects the x-side of the cell (x or x+1), each lane
// walks the 4 (y,z) corners of its side.  The hash uses prime 1 for x, so the two x-neighbours are adjacent
// table entries: a warp-wide gather / reduction instruction then touches at most 16 cache lines instead of
// 32 -- L1TEX line processing (~2 cycles per line per instruction) is what bounds these kernels, not HBM.
// Points are ordered [ray, sample]; consecutive samples of a ray share cells on coarse levels, so the
// backward merges equal-cell runs inside the warp with a segmented shuffle reduction before issuing
// red.global.add.v2.f32 (warp-aggregated atomics).
#include "lb_common.cuh"

namespace {

constexpr int kThreads = 128;
constexpr int kPointsPerBlock = kThreads / 2;
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
__device__ __noinline__ uint4 wrap_indices(uint4 i, uint32_t size) {
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

// ------------------------------------------------------------------------------------------
// forward: enc[p][2l+f] = sum over 8 corners of w_c * table[idx_c][f]
// ------------------------------------------------------------------------------------------
template <bool kFromRays, int kNDense>
__global__ void __launch_bounds__(kThreads, 8)
encode_fwd_kernel(lb_level_table T, const float* __restrict__ rays, const float* __restrict__ zv,
                  const float* __restrict__ x01, const float2* __restrict__ table, size_t n_points, int n_samples,
                  float* __restrict__ enc, int l_begin, int l_end) {
    const int half = threadIdx.x & 1;
    const int width = T.n_levels * LB_FEATS;
    const size_t n_tiles = (n_points + kPointsPerBlock - 1) / kPointsPerBlock;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t p = tile * kPointsPerBlock + (threadIdx.x >> 1);
        const bool valid = p < n_points;
        float part[LB_MLP_IN];
        Point q = {0.5f, 0.5f, 0.5f};
        if (valid) q = load_point<kFromRays>(rays, zv, x01, p, n_samples, nullptr);
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
        // pair exchange: the even lane finishes features 0-15, the odd lane features 16-31
        float out[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float send = half ? part[i] : part[16 + i];
            const float recv = __shfl_xor_sync(kFull, send, 1);
            out[i] = (half ? part[16 + i] : part[i]) + recv;
        }
        if (valid) {
            if (width == LB_MLP_IN) {
                // one float4 = two levels; a pass stores only its own columns (passes start on even levels)
                float4* dst = reinterpret_cast<float4*>(enc + p * LB_MLP_IN + half * 16);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int l0 = 2 * (half * 4 + i);
                    if (l0 >= l_begin && l0 < l_end) __stcs(dst + i, make_float4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]));
                }
            } else {
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int c = half * 16 + i;
                    if (c < width && c >= 2 * l_begin && c < 2 * l_end) enc[p * width + c] = out[i];
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// backward: table scatter-add (warp-aggregated) and optional dL/dx01
// ------------------------------------------------------------------------------------------
constexpr int kAggMaxRuns = 12;   // aggregate when the warp's 16 points fall into <= 12 distinct cells
constexpr int kAggLevels = 8;     // only the coarse levels carry the run-merging code (instruction-cache budget);
                                  // finer levels always scatter every corner, which is equally correct

// segmented suffix sum over the warp's 16 points (2 lanes each): a run head collects its whole run.  Out of
// line: one copy instead of 8 values x 8 levels, called convergently by all 32 lanes.
__device__ __noinline__ float seg_suffix_sum(float v, int dist) {
#pragma unroll
    for (int off = 1; off < 16; off <<= 1) {
        const float t = __shfl_down_sync(kFull, v, 2 * off);
        if (off < dist) v += t;
    }
    return v;
}

template <bool kFromRays, bool kInputGrad, int kNDense>
__global__ void __launch_bounds__(kThreads, 6)
encode_bwd_kernel(lb_level_table T, const float* __restrict__ rays, const float* __restrict__ zv,
                  const float* __restrict__ x01, const float2* __restrict__ table, const float* __restrict__ d_enc,
                  size_t n_points, int n_samples, float* __restrict__ grad_table, float* __restrict__ d_rays,
                  float* __restrict__ d_x01, int l_begin, int l_end) {
    const int lane = threadIdx.x & 31;
    const int half = lane & 1;
    const int pair = lane >> 1;
    const int width = T.n_levels * LB_FEATS;
    const size_t n_tiles = (n_points + kPointsPerBlock - 1) / kPointsPerBlock;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t p = tile * kPointsPerBlock + (threadIdx.x >> 1);
        const bool valid = p < n_points;
        float de[LB_MLP_IN];
        Point q = {0.5f, 0.5f, 0.5f};
        float zval = 0.0f;
        if (valid) {
            q = load_point<kFromRays>(rays, zv, x01, p, n_samples, &zval);
            if (width == LB_MLP_IN) {
                const float4* src = reinterpret_cast<const float4*>(d_enc + p * LB_MLP_IN);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (2 * i >= l_begin && 2 * i < l_end) v = __ldcs(src + i);    // only this pass's levels; use-once stream
                    de[4 * i] = v.x; de[4 * i + 1] = v.y; de[4 * i + 2] = v.z; de[4 * i + 3] = v.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < LB_MLP_IN; ++i)
                    de[i] = (i < width && i >= 2 * l_begin && i < 2 * l_end) ? __ldg(d_enc + p * width + i) : 0.0f;
            }
        } else {
#pragma unroll
            for (int i = 0; i < LB_MLP_IN; ++i) de[i] = 0.0f;
        }
        // A sample whose upstream gradient row is exactly zero (ReLU-inactive density: dL/dsigma = 0, roughly
        // every second sample while the field is young) scatters nothing: the L2 atomics are what bounds this
        // kernel, so such points issue none.
        uint32_t nz_bits = 0u;
#pragma unroll
        for (int i = 0; i < LB_MLP_IN; ++i) nz_bits |= __float_as_uint(de[i]);
        const bool nonzero = (nz_bits << 1) != 0u;
        if (__all_sync(kFull, !nonzero)) {                             // warp-uniform: nothing to scatter for these 16 points
            if (kInputGrad && !kFromRays && valid && !half && l_begin == 0) { d_x01[3 * p + 0] = 0.0f; d_x01[3 * p + 1] = 0.0f; d_x01[3 * p + 2] = 0.0f; }
            continue;
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
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            a[k] = seg_suffix_sum(a[k], dist);
                            b[k] = seg_suffix_sum(b[k], dist);
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
}

int grid_for(int sm_count, size_t n_points, int blocks_per_sm) {
    const size_t tiles = (n_points + kPointsPerBlock - 1) / kPointsPerBlock;
    const size_t cap = (size_t)sm_count * blocks_per_sm;
    return (int)(tiles < cap ? (tiles ? tiles : 1) : cap);
}

// number of leading dense levels if the plan is "dense prefix, hashed (power-of-two) rest", else -1
int dense_prefix(const lb_plan* plan) {
    int nd = 0;
    while (nd < plan->cfg.n_levels && plan->levels[nd].dense) ++nd;
    for (int l = nd; l < plan->cfg.n_levels; ++l)
        if (plan->levels[l].dense || plan->levels[l].mask == 0u) return -1;
    return (nd >= 1 && nd <= 4) ? nd : -1;
}

template <int kNDense>
void launch_fwd(const lb_level_table& T, int grid, const float* rays, const float* z, const float* x01, const float2* tb,
                size_t n_points, int n_samples, float* enc, int l0, int l1, cudaStream_t st) {
    if (rays) encode_fwd_kernel<true, kNDense><<<grid, kThreads, 0, st>>>(T, rays, z, nullptr, tb, n_points, n_samples, enc, l0, l1);
    else encode_fwd_kernel<false, kNDense><<<grid, kThreads, 0, st>>>(T, nullptr, nullptr, x01, tb, n_points, 1, enc, l0, l1);
}

template <int kNDense>
void launch_bwd(const lb_level_table& T, int grid, const float* rays, const float* z, const float* x01, const float2* tb,
                const float* d_enc, size_t n_points, int n_samples, float* grad_table, float* d_rays, float* d_x01,
                int l0, int l1, cudaStream_t st) {
    const bool in_grad = (d_rays != nullptr) || (d_x01 != nullptr);
    if (rays) {
        if (in_grad) encode_bwd_kernel<true, true, kNDense><<<grid, kThreads, 0, st>>>(T, rays, z, nullptr, tb, d_enc, n_points, n_samples, grad_table, d_rays, nullptr, l0, l1);
        else encode_bwd_kernel<true, false, kNDense><<<grid, kThreads, 0, st>>>(T, rays, z, nullptr, tb, d_enc, n_points, n_samples, grad_table, nullptr, nullptr, l0, l1);
    } else {
        if (in_grad) encode_bwd_kernel<false, true, kNDense><<<grid, kThreads, 0, st>>>(T, nullptr, nullptr, x01, tb, d_enc, n_points, 1, grad_table, nullptr, d_x01, l0, l1);
        else encode_bwd_kernel<false, false, kNDense><<<grid, kThreads, 0, st>>>(T, nullptr, nullptr, x01, tb, d_enc, n_points, 1, grad_table, nullptr, nullptr, l0, l1);
    }
}

// Level-group passes.  When the tables fit the L2 (T <= 2^19: 57 MB) one launch walks all levels per point.  Larger
// tables (T = 2^22: 422 MB) would turn every gather / atomic into a random 32-byte HBM access, so the levels are
// processed in groups whose tables stay L2-resident while ALL points of the chunk visit them: one launch per
// group, each reading / writing only its own columns of the row-major [P, 32] encoding.  Groups start on even
// levels (a float4 of the row = two levels).  Returns the end of the group that starts at l0.
int next_group_end(const lb_plan* plan, int l0) {
    const int n = plan->cfg.n_levels;
    const uint64_t budget = (uint64_t)plan->l2_bytes * 55 / 100;
    if ((uint64_t)plan->n_table_entries * 8 <= budget) return n;
    uint64_t bytes = 0;
    int l = l0;
    while (l < n) {
        uint64_t step = (uint64_t)plan->levels[l].size * 8;
        if (l + 1 < n) step += (uint64_t)plan->levels[l + 1].size * 8;
        if (l > l0 && bytes + step > budget) break;
        bytes += step;
        l += 2;
    }
    return l < n ? l : n;
}

}  // namespace

// internal launchers shared with lb_field.cu
int lb_launch_encode_fwd(const lb_plan* plan, const float* rays, const float* z, const float* x01, const float* table,
                         size_t n_points, int n_samples, float* enc, cudaStream_t st) {
    const lb_level_table T = lb_make_level_table(plan);
    const int grid = grid_for(plan->sm_count, n_points, 8);
    const float2* tb = reinterpret_cast<const float2*>(table);
    const int nd = dense_prefix(plan);
    for (int l0 = 0; l0 < plan->cfg.n_levels;) {
        const int l1 = next_group_end(plan, l0);
        switch (nd) {
            case 1: launch_fwd<1>(T, grid, rays, z, x01, tb, n_points, n_samples, enc, l0, l1, st); break;
            case 2: launch_fwd<2>(T, grid, rays, z, x01, tb, n_points, n_samples, enc, l0, l1, st); break;
            case 3: launch_fwd<3>(T, grid, rays, z, x01, tb, n_points, n_samples, enc, l0, l1, st); break;
            case 4: launch_fwd<4>(T, grid, rays, z, x01, tb, n_points, n_samples, enc, l0, l1, st); break;
            default: launch_fwd<-1>(T, grid, rays, z, x01, tb, n_points, n_samples, enc, l0, l1, st); break;
        }
        l0 = l1;
    }
    return lb_check_cuda(cudaGetLastError(), "encode_fwd launch");
}

int lb_launch_encode_bwd(const lb_plan* plan, const float* rays, const float* z, const float* x01, const float* table,
                         const float* d_enc, size_t n_points, int n_samples, float* grad_table, float* d_rays,
                         float* d_x01, cudaStream_t st) {
    const lb_level_table T = lb_make_level_table(plan);
    const int grid = grid_for(plan->sm_count, n_points, 6);
    const float2* tb = reinterpret_cast<const float2*>(table);
    const int nd = dense_prefix(plan);
    for (int l0 = 0; l0 < plan->cfg.n_levels;) {
        const int l1 = next_group_end(plan, l0);
        switch (nd) {
            case 1: launch_bwd<1>(T, grid, rays, z, x01, tb, d_enc, n_points, n_samples, grad_table, d_rays, d_x01, l0, l1, st); break;
            case 2: launch_bwd<2>(T, grid, rays, z, x01, tb, d_enc, n_points, n_samples, grad_table, d_rays, d_x01, l0, l1, st); break;
            case 3: launch_bwd<3>(T, grid, rays, z, x01, tb, d_enc, n_points, n_samples, grad_table, d_rays, d_x01, l0, l1, st); break;
            case 4: launch_bwd<4>(T, grid, rays, z, x01, tb, d_enc, n_points, n_samples, grad_table, d_rays, d_x01, l0, l1, st); break;
            default: launch_bwd<-1>(T, grid, rays, z, x01, tb, d_enc, n_points, n_samples, grad_table, d_rays, d_x01, l0, l1, st); break;
        }
        l0 = l1;
    }
    return lb_check_cuda(cudaGetLastError(), "encode_bwd launch");
}

extern "C" int lb_encode_fwd(const lb_plan* plan, const float* x01, const float* table, int64_t n_points,
                             float* enc, void* stream) {
    LB_REQUIRE(plan && x01 && table && enc, "null pointer");
    LB_REQUIRE(n_points >= 0 && n_points < (int64_t)1 << 32, "n_points must be in [0, 2^32)");
    if (n_points == 0) return LB_OK;
    return lb_launch_encode_fwd(plan, nullptr, nullptr, x01, table, (size_t)n_points, 1, enc, (cudaStream_t)stream);
}

extern "C" int lb_encode_bwd(const lb_plan* plan, const float* x01, const float* table, const float* d_enc,
                             int64_t n_points, float* grad_table, float* d_x01, void* stream) {
    LB_REQUIRE(plan && x01 && table && d_enc, "null pointer");
    LB_REQUIRE(grad_table || d_x01, "nothing to compute");
    LB_REQUIRE(n_points >= 0 && n_points < (int64_t)1 << 32, "n_points must be in [0, 2^32)");
    if (n_points == 0) return LB_OK;
    return lb_launch_encode_bwd(plan, nullptr, nullptr, x01, table, d_enc, (size_t)n_points, 1, grad_table, nullptr,
                                d_x01, (cudaStream_t)stream);
}
