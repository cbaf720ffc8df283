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
#include "lb_encode.cuh"

namespace {

using namespace lb_enc;

constexpr int kThreads = 128;
constexpr int kPointsPerBlock = kThreads / 2;

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
        Point q = {0.5f, 0.5f, 0.5f};
        if (valid) q = load_point<kFromRays>(rays, zv, x01, p, n_samples, nullptr);
        float out[16];
        encode_fwd_point<kNDense>(T, q, half, table, l_begin, l_end, out);
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
// level-major hand-off (tables beyond the L2: T >= 2^20).  The encode <-> MLP hand-off of lb_sigma_fwd / lb_sigma_bwd is
// then stored as 16 planes planes[l][p] (float2) instead of rows [p][32]: a level-group pass streams ONLY its own planes
// (8 B per point and level, whole sectors), where a pass over row-major rows touches 8 B of every 128 B row -- a
// 32-byte sector per point for the forward's column stores, the whole 1.07 GB of gradient rows per backward pass.  That
// is what makes single-level passes (one 33.5 MB table resident in the L2 while all points visit it) pay.
// Lean forward kernel: no 16-way unrolled level loop, one pair exchange per level, bit-identical to encode_fwd_point.
// ------------------------------------------------------------------------------------------
template <bool kFromRays>
__global__ void __launch_bounds__(kThreads, 16)
encode_fwd_lm_kernel(lb_level_table T, const float* __restrict__ rays, const float* __restrict__ zv,
                     const float* __restrict__ x01, const float2* __restrict__ table, size_t n_points, int n_samples,
                     float2* __restrict__ planes, int l_begin, int l_end) {
    const int half = threadIdx.x & 1;
    const size_t n_tiles = (n_points + kPointsPerBlock - 1) / kPointsPerBlock;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t p = tile * kPointsPerBlock + (threadIdx.x >> 1);
        const bool valid = p < n_points;
        Point q = {0.5f, 0.5f, 0.5f};
        if (valid) q = load_point<kFromRays>(rays, zv, x01, p, n_samples, nullptr);
        for (int l = l_begin; l < l_end; ++l) {
            const lb_level& L = T.lv[l];
            const Side s = L.dense ? locate_side_t<true>(L, q, half) : locate_side_t<false>(L, q, half);
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
            a *= s.wx;
            b *= s.wx;
            a += __shfl_xor_sync(kFull, a, 1);
            b += __shfl_xor_sync(kFull, b, 1);
            if (valid && !half) __stcs(planes + (size_t)l * n_points + p, make_float2(a, b));
        }
    }
}

// Lean backward of a level group, table gradient only (dL/dx passes use encode_bwd_kernel): runtime level loop, the
// pass's own planes only, 12 blocks per SM -- a single-level pass is two dependent memory round trips (z / plane, then
// the reductions) per warp, so residency, not instruction count, sets its speed (the 16-level general kernel at 6 blocks
// per SM took 0.40 ms per single-level pass of 8.4 M points, this one 0.29 ms; 16 blocks per SM changed nothing: the L2
// reduction rate binds).  Same run merging on the coarse levels
// as lb_enc::encode_bwd_point, with its own thresholds (below).
#ifndef LB_BWD_LM_BLOCKS
#define LB_BWD_LM_BLOCKS 12
#endif
// Run merging (lb_encode.cuh) in the level-major passes: levels below LB_AGG_LEVELS_LM merge runs of equal cells before
// the reductions when the warp's 16 points fall into <= LB_AGG_MAX_RUNS_LM cells.  Measured inside the c4 bench step
// (T = 2^22, lb_sigma_bwd per step, profiles/run_variants.sh): (levels, runs) = (0, -) 21.56 ms -- merging is what makes
// the coarse levels affordable -- (4, 15) 8.41, (6, 15) 8.17, (8, 15) 8.24, (12, 15) 8.31, (8, 8) 8.15.
#ifndef LB_AGG_LEVELS_LM
#define LB_AGG_LEVELS_LM 8
#define LB_AGG_MAX_RUNS_LM 8
#endif
template <bool kFromRays>
__global__ void __launch_bounds__(kThreads, LB_BWD_LM_BLOCKS)
encode_bwd_lm_kernel(lb_level_table T, const float* __restrict__ rays, const float* __restrict__ zv,
                     const float* __restrict__ x01, const float2* __restrict__ planes, size_t n_points, int n_samples,
                     float* __restrict__ grad_table, int l_begin, int l_end) {
    const int lane = threadIdx.x & 31;
    const int half = lane & 1;
    const int pair = lane >> 1;
    const size_t n_tiles = (n_points + kPointsPerBlock - 1) / kPointsPerBlock;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t p = tile * kPointsPerBlock + (threadIdx.x >> 1);
        const bool valid = p < n_points;
        Point q = {0.5f, 0.5f, 0.5f};
        if (valid) q = load_point<kFromRays>(rays, zv, x01, p, n_samples, nullptr);
        bool all_distinct = false;
        for (int l = l_begin; l < l_end; ++l) {
            float2 d = make_float2(0.f, 0.f);
            if (valid) d = __ldcs(planes + (size_t)l * n_points + p);
            const bool nonzero = ((__float_as_uint(d.x) | __float_as_uint(d.y)) << 1) != 0u;
            if (__all_sync(kFull, !nonzero)) continue;               // ReLU-inactive samples scatter nothing
            const lb_level& L = T.lv[l];
            const Side s = L.dense ? locate_side_t<true>(L, q, half) : locate_side_t<false>(L, q, half);
            float a[4], b[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float w = s.wx * s.wy[k & 1] * s.wz[k >> 1];
                a[k] = w * d.x;
                b[k] = w * d.y;
            }
            bool issue = valid && nonzero;
            if (l < LB_AGG_LEVELS_LM && !all_distinct) {
                const uint32_t pcx = __shfl_up_sync(kFull, s.cx, 2), pcy = __shfl_up_sync(kFull, s.cy, 2),
                               pcz = __shfl_up_sync(kFull, s.cz, 2);
                const int pvalid = __shfl_up_sync(kFull, (int)valid, 2);
                const bool same = (lane >= 2) && valid && pvalid && pcx == s.cx && pcy == s.cy && pcz == s.cz;
                const unsigned heads = __ballot_sync(kFull, !same) & 0x55555555u;
                const int runs = __popc(heads);
                if (runs == 16) {
                    all_distinct = true;
                } else if (runs <= LB_AGG_MAX_RUNS_LM) {
                    const unsigned rest = (pair < 15) ? (heads >> (2 * pair + 2)) : 0u;
                    const int dist = rest ? 1 + ((__ffs(rest) - 1) >> 1) : 16 - pair;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        a[k] = seg_suffix_sum(a[k], dist);
                        b[k] = seg_suffix_sum(b[k], dist);
                    }
                    issue = valid && !same;
                }
            }
            if (issue) {
#pragma unroll
                for (int k = 0; k < 4; ++k) red_add_f32x2(grad_table + 2 * (size_t)s.idx[k], a[k], b[k]);
            }
        }
    }
}

template <bool kFromRays, bool kInputGrad, int kNDense>
__global__ void __launch_bounds__(kThreads, 6)
encode_bwd_kernel(lb_level_table T, const float* __restrict__ rays, const float* __restrict__ zv,
                  const float* __restrict__ x01, const float2* __restrict__ table, const float* __restrict__ d_enc,
                  size_t n_points, int n_samples, float* __restrict__ grad_table, float* __restrict__ d_rays,
                  float* __restrict__ d_x01, int l_begin, int l_end, int level_major) {
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
            if (level_major) {             // planes[l][p] (float2): this pass's planes only, both lanes of the pair
                const float2* planes = reinterpret_cast<const float2*>(d_enc);
#pragma unroll
                for (int l = 0; l < 16; ++l) {
                    float2 v = make_float2(0.f, 0.f);
                    if (l >= l_begin && l < l_end) v = __ldcs(planes + (size_t)l * n_points + p);
                    de[2 * l] = v.x; de[2 * l + 1] = v.y;
                }
            } else if (width == LB_MLP_IN) {
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
        lb_enc::encode_bwd_point<kFromRays, kInputGrad, kNDense>(T, q, zval, de, valid, p, n_samples, table, grad_table,
                                                                 d_rays, d_x01, l_begin, l_end);
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
                int l0, int l1, int lm, cudaStream_t st) {
    const bool in_grad = (d_rays != nullptr) || (d_x01 != nullptr);
    if (rays) {
        if (in_grad) encode_bwd_kernel<true, true, kNDense><<<grid, kThreads, 0, st>>>(T, rays, z, nullptr, tb, d_enc, n_points, n_samples, grad_table, d_rays, nullptr, l0, l1, lm);
        else encode_bwd_kernel<true, false, kNDense><<<grid, kThreads, 0, st>>>(T, rays, z, nullptr, tb, d_enc, n_points, n_samples, grad_table, nullptr, nullptr, l0, l1, lm);
    } else {
        if (in_grad) encode_bwd_kernel<false, true, kNDense><<<grid, kThreads, 0, st>>>(T, nullptr, nullptr, x01, tb, d_enc, n_points, 1, grad_table, nullptr, d_x01, l0, l1, lm);
        else encode_bwd_kernel<false, false, kNDense><<<grid, kThreads, 0, st>>>(T, nullptr, nullptr, x01, tb, d_enc, n_points, 1, grad_table, nullptr, nullptr, l0, l1, lm);
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

// Groups of the level-major path: levels are taken greedily while their tables fit 38 % of the L2 (48 MB of 126: a
// randomly addressed table stays resident only while it fits one of the two L2 partitions next to the streams of the
// pass -- profiles/README.md, "T = 2^22"); a level that does not fit by itself forms its own group.  T = 2^22: the dense
// levels 0-3 (19.7 MB), then twelve single hashed levels of 33.5 MB.  No even-level constraint (planes, not float4 rows).
int next_group_end_lm(const lb_plan* plan, int l0) {
    const int n = plan->cfg.n_levels;
    const uint64_t budget = (uint64_t)plan->l2_bytes * 38 / 100;
    uint64_t bytes = 0;
    int l = l0;
    while (l < n) {
        const uint64_t step = (uint64_t)plan->levels[l].size * 8;
        if (l > l0 && bytes + step > budget) break;
        bytes += step;
        ++l;
    }
    return l;
}

}  // namespace

int lb_encode_dense_prefix(const lb_plan* plan) { return dense_prefix(plan); }
bool lb_encode_single_group(const lb_plan* plan) { return next_group_end(plan, 0) >= plan->cfg.n_levels; }

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
                         float* d_x01, cudaStream_t st, bool level_major) {
    const lb_level_table T = lb_make_level_table(plan);
    const int grid = grid_for(plan->sm_count, n_points, 6);
    const float2* tb = reinterpret_cast<const float2*>(table);
    const int nd = dense_prefix(plan);
    const int lm = level_major ? 1 : 0;
    if (level_major && grad_table && !d_rays && !d_x01) {       // table gradient only: lean single-group passes
        const int grid12 = grid_for(plan->sm_count, n_points, LB_BWD_LM_BLOCKS);
        const float2* pl = reinterpret_cast<const float2*>(d_enc);
        for (int l0 = 0; l0 < plan->cfg.n_levels;) {
            const int l1 = next_group_end_lm(plan, l0);
            if (rays) encode_bwd_lm_kernel<true><<<grid12, kThreads, 0, st>>>(T, rays, z, nullptr, pl, n_points, n_samples, grad_table, l0, l1);
            else encode_bwd_lm_kernel<false><<<grid12, kThreads, 0, st>>>(T, nullptr, nullptr, x01, pl, n_points, 1, grad_table, l0, l1);
            l0 = l1;
        }
        return lb_check_cuda(cudaGetLastError(), "encode_bwd (level-major) launch");
    }
    for (int l0 = 0; l0 < plan->cfg.n_levels;) {
        const int l1 = level_major ? next_group_end_lm(plan, l0) : next_group_end(plan, l0);
        switch (nd) {
            case 1: launch_bwd<1>(T, grid, rays, z, x01, tb, d_enc, n_points, n_samples, grad_table, d_rays, d_x01, l0, l1, lm, st); break;
            case 2: launch_bwd<2>(T, grid, rays, z, x01, tb, d_enc, n_points, n_samples, grad_table, d_rays, d_x01, l0, l1, lm, st); break;
            case 3: launch_bwd<3>(T, grid, rays, z, x01, tb, d_enc, n_points, n_samples, grad_table, d_rays, d_x01, l0, l1, lm, st); break;
            case 4: launch_bwd<4>(T, grid, rays, z, x01, tb, d_enc, n_points, n_samples, grad_table, d_rays, d_x01, l0, l1, lm, st); break;
            default: launch_bwd<-1>(T, grid, rays, z, x01, tb, d_enc, n_points, n_samples, grad_table, d_rays, d_x01, l0, l1, lm, st); break;
        }
        l0 = l1;
    }
    return lb_check_cuda(cudaGetLastError(), "encode_bwd launch");
}

// forward into level-major planes [n_levels][n_points] (float2): one lean launch per L2-resident level group
int lb_launch_encode_fwd_lm(const lb_plan* plan, const float* rays, const float* z, const float* x01, const float* table,
                            size_t n_points, int n_samples, float* planes, cudaStream_t st) {
    const lb_level_table T = lb_make_level_table(plan);
    const int grid = grid_for(plan->sm_count, n_points, 16);       // 32 registers: 2048 threads per SM
    const float2* tb = reinterpret_cast<const float2*>(table);
    float2* pl = reinterpret_cast<float2*>(planes);
    for (int l0 = 0; l0 < plan->cfg.n_levels;) {
        const int l1 = next_group_end_lm(plan, l0);
        if (rays) encode_fwd_lm_kernel<true><<<grid, kThreads, 0, st>>>(T, rays, z, nullptr, tb, n_points, n_samples, pl, l0, l1);
        else encode_fwd_lm_kernel<false><<<grid, kThreads, 0, st>>>(T, nullptr, nullptr, x01, tb, n_points, 1, pl, l0, l1);
        l0 = l1;
    }
    return lb_check_cuda(cudaGetLastError(), "encode_fwd (level-major) launch");
}

extern "C" int lb_encode_fwd(const lb_plan* plan, const float* x01, const float* table, int64_t n_points,
                             float* enc, void* stream) {
    LB_REQUIRE(plan, "null pointer");
    LB_REQUIRE(n_points == 0 || (x01 && table && enc), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_points >= 0 && n_points < (int64_t)1 << 32, "n_points must be in [0, 2^32)");
    if (n_points == 0) return LB_OK;
    return lb_launch_encode_fwd(plan, nullptr, nullptr, x01, table, (size_t)n_points, 1, enc, (cudaStream_t)stream);
}

extern "C" int lb_encode_bwd(const lb_plan* plan, const float* x01, const float* table, const float* d_enc,
                             int64_t n_points, float* grad_table, float* d_x01, void* stream) {
    LB_REQUIRE(plan, "null pointer");
    LB_REQUIRE(n_points == 0 || (x01 && table && d_enc), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(grad_table || d_x01, "nothing to compute");
    LB_REQUIRE(n_points >= 0 && n_points < (int64_t)1 << 32, "n_points must be in [0, 2^32)");
    if (n_points == 0) return LB_OK;
    return lb_launch_encode_bwd(plan, nullptr, nullptr, x01, table, d_enc, (size_t)n_points, 1, grad_table, nullptr,
                                d_x01, (cudaStream_t)stream, false);
}
