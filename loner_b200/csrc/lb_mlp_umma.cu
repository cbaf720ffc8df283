// R4 (network half) on the 5th-generation tensor cores: the 32 -> 64 -> 1 sigma MLP, forward and backward,
// as tcgen05.mma (UMMA) kernels with fp32 accumulators in tensor memory (TMEM).
//
// Replaces tcnn FullyFusedMLP inside tcnn.NetworkWithInputEncoding (reference src/models/nerf_tcnn.py:35-38,71;
// cfg/nerf_config/default_nerf_hash.yaml:26-31).  Weight layout = tcnn's: W1 [64,32] then W2 [16,64] row-major.
//
// Precision.  north_star wants 1e-3 agreement with fp32 accumulate, and the weight gradient is a strongly
// cancelling sum, so a single bf16/tf32 pass is not enough.  Every fp32 operand x is split into bf16 pieces
// x = b1 + b2 (+ O(2^-18 x)), b1 = rn_bf16(x), b2 = rn_bf16(x - b1), and the products are accumulated in fp32:
//   forward : kind::tf32 on exact hi/lo splits of the fp32 operands, H = Ehi Whi^T + Ehi Wlo^T + Elo Whi^T
//             (error ~2^-21: the ReLU mask must be as reliable as an fp32 evaluation, a flipped unit changes a
//             gradient row by O(1); the split costs 2 ALU ops per element and needs no lane exchange)
//   backward: the A operand is the ReLU mask m[p,j] in {0,1}, exact in bf16, so two pieces of B suffice:
//     dE[p,:] = ds[p] * sum_j m[p,j] (w2[j] W1[j,:])                  A = M (K-major),  B = W1s pieces
//     U[j,k]  = sum_p m[p,j] (ds[p] e[p,k])                           A = M^T (MN-major view of the same tile)
//     dW1[j,k] = w2[j] U[j,k],  dW2[j] = sum_k W1[j,k] U[j,k]         (epilogue, fp32)
//   The mask is produced by the forward kernel (64 bits per point) so the backward needs no recomputation.
//
// Shared-memory operand layout (no swizzle, "interleaved" canonical UMMA layout): a tile X[r][c] of 16-byte
// chunks (8 bf16) is stored at  (r % 8) * 16 + (r / 8) * 128 + c * LBO  with LBO = (rows * 16 + 16) bytes.  The
// extra 16 bytes rotate consecutive chunks over the banks so the conversion stores are conflict free.  The same
// bytes serve as a K-major operand (rows = M/N, chunks along K: desc.SBO = 128, desc.LBO = LBO) and as an
// MN-major operand (rows = K, chunks along M/N: desc.SBO = LBO, desc.LBO = 128).
//
// One CTA = 128 threads = one 128-point tile at a time (UMMA M = 128), persistent over tiles, 3 CTAs per SM
// so that one CTA's conversion / epilogue overlaps another's MMAs and global loads.
#include <cuda_bf16.h>

#include <stdlib.h>

#include "lb_common.cuh"
#include "lb_encode.cuh"
#include "lb_umma.cuh"

namespace {

using namespace lb_umma;

constexpr int kTile = 128;
constexpr int kThreads = 128;
constexpr uint32_t kLboP = kTile * 16 + 16;        // chunk stride of a 128-row tile (points)
constexpr uint32_t kLboW = LB_MLP_W * 16 + 16;     // chunk stride of a 64-row tile (hidden units)
constexpr uint32_t kTmemCols = 64;
// resident CTAs per SM (shared memory 49.7 KB forward / 42.5 KB backward, 64 TMEM columns each: 4 x 64 <= 512 columns)
#ifndef LB_MLP_FWD_CTAS
#define LB_MLP_FWD_CTAS 4
#endif
#ifndef LB_MLP_BWD_CTAS      // c4 bench step (level-major planes, T = 2^22): 3 / 4 / 5 CTAs per SM = 14.10 / 14.04 / 14.41 ms
#define LB_MLP_BWD_CTAS 4
#endif
constexpr int kFwdCtasPerSm = LB_MLP_FWD_CTAS;
constexpr int kBwdCtasPerSm = LB_MLP_BWD_CTAS;


// A 128 x 32 fp32 tile (row-major in global memory, 8 float4 per row) is handled "flat": in pass i thread t owns
// float4 number q = t % 8 of row i * 16 + t / 8, so every warp-wide global access is 512 contiguous bytes.
// Lane pairs (q even / odd) exchange their float4 to assemble one 8-element chunk; the even lane then stores the
// chunk of first pieces, the odd lane the chunk of second pieces.
struct TileRegs {
    float4 f[8];
};

__device__ __forceinline__ void load_tile(const float* __restrict__ src, size_t tile, size_t n_points, TileRegs& t) {
    const float4* base = reinterpret_cast<const float4*>(src) + tile * (size_t)(kTile * 8);
    const size_t first = tile * (size_t)kTile;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = i * 16 + (threadIdx.x >> 3);
        t.f[i] = (first + row < n_points) ? __ldg(base + i * kThreads + threadIdx.x) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// the same register tile from level-major planes planes[l][p] (float2): float4 q of a row = levels 2q and 2q + 1
__device__ __forceinline__ void load_tile_lm(const float* __restrict__ src, size_t tile, size_t n_points, TileRegs& t) {
    const float2* planes = reinterpret_cast<const float2*>(src);
    const size_t first = tile * (size_t)kTile;
    const int q = threadIdx.x & 7;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const size_t p = first + i * 16 + (threadIdx.x >> 3);
        if (p < n_points) {
            const float2 a = __ldg(planes + (size_t)(2 * q) * n_points + p);
            const float2 b = __ldg(planes + (size_t)(2 * q + 1) * n_points + p);
            t.f[i] = make_float4(a.x, a.y, b.x, b.y);
        } else {
            t.f[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
}

// converts the register tile (optionally scaled per row by s_scale[row]) into two or three bf16 piece tiles
template <bool kThree>
__device__ __forceinline__ void convert_tile(const TileRegs& t, const float* s_scale, uint8_t* p1, uint8_t* p2,
                                             uint8_t* p3) {
    const int odd = threadIdx.x & 1;
    const int chunk = (threadIdx.x & 7) >> 1;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = i * 16 + (threadIdx.x >> 3);
        float4 mine = t.f[i];
        if (s_scale) {
            const float s = s_scale[row];
            mine.x *= s; mine.y *= s; mine.z *= s; mine.w *= s;
        }
        float4 other;
        other.x = __shfl_xor_sync(0xffffffffu, mine.x, 1);
        other.y = __shfl_xor_sync(0xffffffffu, mine.y, 1);
        other.z = __shfl_xor_sync(0xffffffffu, mine.z, 1);
        other.w = __shfl_xor_sync(0xffffffffu, mine.w, 1);
        const float4 a = odd ? other : mine, b = odd ? mine : other;       // a = k 0..3 of the chunk, b = k 4..7
        const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        uint32_t b1[4], b2[4], b3[4];
        split8<kThree>(v, b1, b2, b3);
        const uint32_t off = tile_off(row, chunk, kLboP);
        *reinterpret_cast<uint4*>((odd ? p2 : p1) + off) =
            odd ? make_uint4(b2[0], b2[1], b2[2], b2[3]) : make_uint4(b1[0], b1[1], b1[2], b1[3]);
        if (kThree && !odd) *reinterpret_cast<uint4*>(p3 + off) = make_uint4(b3[0], b3[1], b3[2], b3[3]);
    }
}

// W[64][32] fp32 (row scale optional) -> two bf16 piece tiles X[r = j][c = k / 8]
__device__ __forceinline__ void convert_weights(const float* __restrict__ w1, const float* __restrict__ row_scale,
                                                uint8_t* p1, uint8_t* p2, uint8_t* p3) {
    for (int it = threadIdx.x; it < LB_MLP_W * 4; it += kThreads) {
        const int j = it >> 2, c = it & 3;
        const float s = row_scale ? __ldg(row_scale + j) : 1.0f;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = __ldg(w1 + j * LB_MLP_IN + c * 8 + e) * s;
        uint32_t b1[4], b2[4], b3[4];
        split8<true>(v, b1, b2, b3);
        *reinterpret_cast<uint4*>(p1 + tile_off(j, c, kLboW)) = make_uint4(b1[0], b1[1], b1[2], b1[3]);
        *reinterpret_cast<uint4*>(p2 + tile_off(j, c, kLboW)) = make_uint4(b2[0], b2[1], b2[2], b2[3]);
        if (p3) *reinterpret_cast<uint4*>(p3 + tile_off(j, c, kLboW)) = make_uint4(b3[0], b3[1], b3[2], b3[3]);
    }
}


__device__ __forceinline__ void convert_weights_tf32(const float* __restrict__ w1, uint8_t* hi_tile, uint8_t* lo_tile) {
    for (int it = threadIdx.x; it < LB_MLP_W * 8; it += kThreads) {
        const int j = it >> 3, c = it & 7;
        const float4 v = __ldg(reinterpret_cast<const float4*>(w1 + j * LB_MLP_IN + c * 4));
        uint4 hi, lo;
        split_tf32(v, hi, lo);
        *reinterpret_cast<uint4*>(hi_tile + tile_off(j, c, kLboW)) = hi;
        *reinterpret_cast<uint4*>(lo_tile + tile_off(j, c, kLboW)) = lo;
    }
}

// ---- shared-memory carve-up -----------------------------------------------------------------
constexpr uint32_t kPieceP = 4 * kLboP;            // [128 x 32] bf16 piece tile
constexpr uint32_t kPieceW = 4 * kLboW;            // [ 64 x 32] bf16 piece tile
constexpr uint32_t kMaskT = 8 * kLboP;             // [128 x 64] bf16 mask tile
constexpr uint32_t kOutStride = 144;               // padded row stride of the fp32 dE staging tile

constexpr uint32_t kRowStride = 144;               // padded row stride of a raw fp32 staging tile (128 B of data)
constexpr uint32_t kStage = kTile * kRowStride;
constexpr uint32_t kFwdTmemCols = 128;             // H accumulator 64 + E hi 32 + E lo 32

struct FwdSmem {      // two raw staging tiles (cp.async ring), tf32 hi / lo weight tiles (8 chunks of 4 elements per row)
    static constexpr uint32_t stage = 0, whi = stage + 2 * kStage, wlo = whi + 8 * kLboW, w2 = wlo + 8 * kLboW,
                              v = w2 + 64 * 4, bar = v + 32 * 4, tmem = bar + 8, total = tmem + 8;
};
struct BwdSmem {      // the fp32 dE staging tile reuses the mask / E' operand tiles, dead once the MMAs have completed
    static constexpr uint32_t mt = 0, e1 = mt + kMaskT, e2 = e1 + kPieceP, wa = e2 + kPieceP, wb = wa + kPieceW,
                              out = 0, ds = wb + kPieceW, bar = ds + kTile * 4, tmem = bar + 8, total = tmem + 8;
    static_assert(kTile * kOutStride <= wa, "dE staging must fit in the operand tiles it aliases");
};

// ------------------------------------------------------------------------------------------
// forward: sigma[p] = sum_j w2[j] relu(sum_k W1[j,k] e[p,k]);  act[p] = bit mask of the active hidden units
//
// The shared-memory data pipe (LSU stores + tensor-core operand reads) was the busiest unit of the first version
// (ncu: 38 % + 44 %), so the A operand (the encodings) now lives in tensor memory: cp.async stages the raw fp32
// tile (coalesced, padded rows), thread t reads row t conflict-free, splits it into tf32 hi / lo and writes both
// to TMEM lane t with tcgen05.st; only the 2 x 8 KB weight tiles are read from shared memory by the MMAs.
// ------------------------------------------------------------------------------------------
// Tile walked by this CTA in round kt of the two fused sigma kernels.  A plain stripe (blockIdx.x + kt * gridDim.x)
// gives every CTA the same segment of its rays in every round whenever gridDim.x is a multiple of the tiles per ray
// (148 SMs, 512 samples = 4 tiles): the occupancy-guided sampler packs half of the samples around the surface, so
// the far segments cost a different time than the near ones and whole CTAs finish early / late (measured: forward
// 1.60 ms on uniform depths, 1.77 ms on surface-concentrated depths).  Rotating the stripe by one CTA per round walks
// every CTA through all segments.  Rounds run to ceil(n_tiles / gridDim.x) for every CTA; a tile >= n_tiles in the
// last round is empty (all points invalid).
__device__ __forceinline__ size_t fused_tile(uint32_t kt) {
    return (size_t)kt * gridDim.x + (blockIdx.x + kt) % gridDim.x;
}
__device__ __forceinline__ uint32_t fused_rounds(size_t n_tiles) {
    return (uint32_t)((n_tiles + gridDim.x - 1) / gridDim.x);
}

__device__ __forceinline__ void stage_tile(const float* __restrict__ enc, size_t tile, size_t n_points, uint32_t dst) {
    const float4* base = reinterpret_cast<const float4*>(enc) + tile * (size_t)(kTile * 8);
    const size_t first = tile * (size_t)kTile;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = i * 16 + (threadIdx.x >> 3);
        const bool ok = first + row < n_points;
        cp_async16(dst + row * kRowStride + (threadIdx.x & 7) * 16, ok ? (const void*)(base + i * kThreads + threadIdx.x) : (const void*)enc,
                   ok ? 16u : 0u);
    }
}

// level-major source: thread t copies the 16 float2 of point `first + t` (one per plane) into row t; a warp reads 256
// contiguous bytes of each plane
__device__ __forceinline__ void stage_tile_lm(const float* __restrict__ enc, size_t tile, size_t n_points, uint32_t dst) {
    const float2* planes = reinterpret_cast<const float2*>(enc);
    const size_t p = tile * (size_t)kTile + threadIdx.x;
    const bool ok = p < n_points;
#pragma unroll
    for (int l = 0; l < 16; ++l)
        cp_async8(dst + threadIdx.x * kRowStride + l * 8, ok ? (const void*)(planes + (size_t)l * n_points + p) : (const void*)enc, ok ? 8u : 0u);
}

__global__ void __launch_bounds__(kThreads, kFwdCtasPerSm)
mlp_fwd_umma_kernel(const float* __restrict__ mlp, const float* __restrict__ enc, size_t n_points,
                    float* __restrict__ sigma, unsigned long long* __restrict__ act, int level_major) {
    extern __shared__ __align__(128) uint8_t sm[];
    const int tid = threadIdx.x, warp = tid >> 5;
    float* w2s = reinterpret_cast<float*>(sm + FwdSmem::w2);
    float* vs = reinterpret_cast<float*>(sm + FwdSmem::v);
    const uint32_t bar = smem_u32(sm + FwdSmem::bar);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + FwdSmem::tmem);
    const uint32_t stage0 = smem_u32(sm + FwdSmem::stage);

    const size_t n_tiles = (n_points + kTile - 1) / kTile;
    size_t tile = blockIdx.x;
    if (tile < n_tiles) {
        if (level_major) stage_tile_lm(enc, tile, n_points, stage0);
        else stage_tile(enc, tile, n_points, stage0);
    }
    cp_async_commit();

    if (warp == 0) tmem_alloc(smem_u32(tmem_slot), kFwdTmemCols);
    if (tid == 32) mbar_init(bar, 1);
    convert_weights_tf32(mlp, sm + FwdSmem::whi, sm + FwdSmem::wlo);
    // relu(h) = (h + |h|) / 2, so  sigma = e . v + sum_j (w2[j] / 2) |h_j|  with  v = W1^T w2 / 2: the linear half is a
    // 32-term fp32 dot product on the staged row, the other half one FFMA per hidden unit (|h| is an operand modifier)
    if (tid < LB_MLP_W) w2s[tid] = 0.5f * __ldg(mlp + LB_MLP_W * LB_MLP_IN + tid);
    if (tid >= 64 && tid < 64 + LB_MLP_IN) {
        const int k = tid - 64;
        float a = 0.0f;
        for (int j = 0; j < LB_MLP_W; ++j)
            a = fmaf(0.5f * __ldg(mlp + LB_MLP_W * LB_MLP_IN + j), __ldg(mlp + j * LB_MLP_IN + k), a);
        vs[k] = a;
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t t_h = tmem, t_ahi = tmem + 64, t_alo = tmem + 96;
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;

    const uint32_t bhi = smem_u32(sm + FwdSmem::whi), blo = smem_u32(sm + FwdSmem::wlo);
    constexpr uint32_t idesc = umma_idesc_tf32(128, LB_MLP_W);

    uint32_t parity = 0, buf = 0;
    float lin = 0.0f;
    for (; tile < n_tiles; tile += gridDim.x, buf ^= 1u) {
        // the other staging tile was read (by its row owners) before the second barrier of the previous iteration
        const size_t next = tile + gridDim.x;
        if (next < n_tiles) {
            if (level_major) stage_tile_lm(enc, next, n_points, stage0 + (buf ^ 1u) * kStage);
            else stage_tile(enc, next, n_points, stage0 + (buf ^ 1u) * kStage);
        }
        cp_async_commit();
        cp_async_wait<1>();
        __syncthreads();                                               // every thread's copies of this tile have landed
        {   // row tid -> tf32 hi / lo -> TMEM lane tid (the previous tile's MMAs were waited on: A is free)
            const uint8_t* row = sm + FwdSmem::stage + buf * kStage + tid * kRowStride;
            lin = 0.0f;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 v = *reinterpret_cast<const float4*>(row + (h * 4 + q) * 16);
                    const float4 c = *reinterpret_cast<const float4*>(vs + (h * 4 + q) * 4);
                    lin = fmaf(v.x, c.x, lin); lin = fmaf(v.y, c.y, lin);
                    lin = fmaf(v.z, c.z, lin); lin = fmaf(v.w, c.w, lin);
                    uint4 a, b;
                    split_tf32(v, a, b);
                    hi[4 * q] = a.x; hi[4 * q + 1] = a.y; hi[4 * q + 2] = a.z; hi[4 * q + 3] = a.w;
                    lo[4 * q] = b.x; lo[4 * q + 1] = b.y; lo[4 * q + 2] = b.z; lo[4 * q + 3] = b.w;
                }
                tmem_st16(t_ahi + lane_off + h * 16, hi);
                tmem_st16(t_alo + lane_off + h * 16, lo);
            }
            tmem_st_wait();
        }
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {                           // K = 32 = 4 x 8 (tf32)
                const uint64_t dbh = umma_desc(bhi + ks * 2 * kLboW, kLboW, 128);
                const uint64_t dbl = umma_desc(blo + ks * 2 * kLboW, kLboW, 128);
                // hi*hi + hi*lo + lo*hi: the hidden pre-activation is as accurate as an fp32 FMA chain (~2^-21), so
                // the ReLU mask flips no more often than any fp32 evaluation would
                umma_tf32_ts(t_h, t_ahi + ks * 8, dbh, idesc, ks > 0);
                umma_tf32_ts(t_h, t_ahi + ks * 8, dbl, idesc, 1);
                umma_tf32_ts(t_h, t_alo + ks * 8, dbh, idesc, 1);
            }
            umma_commit(bar);
        }
        mbar_wait(bar, parity);
        parity ^= 1;
        tc_fence_after();
        // epilogue: thread t owns point tile*128 + t = TMEM lane t.  Everything per hidden unit runs on the FMA pipe:
        // s += (w2/2) |h|;  on = sat(h * 2^126) in {0, 1} (h > 0, normal);  the mask is summed as a float, 16 bits at a time
        float s = lin;
        unsigned long long bits = 0ull;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            uint32_t r[32];
            tmem_ld32(t_h + lane_off + half * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                float mf = 0.0f;
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float h = __uint_as_float(r[g * 16 + j]);
                    s = fmaf(w2s[half * 32 + g * 16 + j], fabsf(h), s);
                    mf = fmaf(__saturatef(h * 8.507059173e37f), (float)(1u << j), mf);
                }
                const uint32_t m16 = __float_as_uint(mf + 8388608.0f) & 0xffffu;
                bits |= (unsigned long long)m16 << (32 * half + 16 * g);
            }
        }
        const size_t p = tile * (size_t)kTile + tid;
        if (p < n_points) {
            sigma[p] = s;
            if (act) act[p] = bits;
        }
    }
    cp_async_wait<0>();
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, kFwdTmemCols);
}

// ------------------------------------------------------------------------------------------
// backward: enc is overwritten in place with dL/d_enc; grad_mlp += (dW1, dW2 row 0)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, kBwdCtasPerSm)
mlp_bwd_umma_kernel(const float* __restrict__ mlp, float* __restrict__ enc, const float* __restrict__ d_sigma,
                    const unsigned long long* __restrict__ act, size_t n_points, float* __restrict__ grad_mlp,
                    int level_major) {
    extern __shared__ __align__(128) uint8_t sm[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float* ds_s = reinterpret_cast<float*>(sm + BwdSmem::ds);
    const uint32_t bar = smem_u32(sm + BwdSmem::bar);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + BwdSmem::tmem);
    const float* w2 = mlp + LB_MLP_W * LB_MLP_IN;

    if (warp == 0) tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    if (tid == 32) mbar_init(bar, 1);
    convert_weights(mlp, w2, sm + BwdSmem::wa, sm + BwdSmem::wb, nullptr);       // W1s[j,k] = w2[j] W1[j,k]
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t t_de = tmem;                 // dE  [128 x 32]  columns  0..31
    const uint32_t t_u = tmem + 32;             // U   [ 64 x 32]  columns 32..63 (M = 64: lanes 0-15 of each warp)

    const uint32_t mt = smem_u32(sm + BwdSmem::mt), e1 = smem_u32(sm + BwdSmem::e1), e2 = smem_u32(sm + BwdSmem::e2);
    const uint32_t wa = smem_u32(sm + BwdSmem::wa), wb = smem_u32(sm + BwdSmem::wb);
    constexpr uint32_t idesc_de = umma_idesc(128, LB_MLP_IN, 0, 1);     // A = M  K-major,  B = W1s MN-major
    constexpr uint32_t idesc_u = umma_idesc(64, LB_MLP_IN, 1, 1);       // A = M^T MN-major, B = E'  MN-major

    const size_t n_tiles = (n_points + kTile - 1) / kTile;
    // two sets of staging registers, prefetch distance 2 (see the forward kernel)
    struct Staged {
        TileRegs regs;
        float ds;
        unsigned long long bits;
    } st_a, st_b;
    auto fetch = [&](Staged& st, size_t t) {
        st.ds = 0.0f;
        st.bits = 0ull;
        if (t < n_tiles) {
            if (level_major) load_tile_lm(enc, t, n_points, st.regs);
            else load_tile(enc, t, n_points, st.regs);
            const size_t p = t * (size_t)kTile + tid;
            if (p < n_points) { st.ds = __ldg(d_sigma + p); st.bits = __ldg(act + p); }
        }
    };
    size_t tile = blockIdx.x;
    fetch(st_a, tile);
    fetch(st_b, tile + gridDim.x);
    uint32_t parity = 0;
    bool first = true;
    auto process = [&](Staged& st) {
        const float ds = st.ds;
        const unsigned long long bits = st.bits;
        ds_s[tid] = ds;
        __syncthreads();                        // also: the previous tile's staging reads are finished
        convert_tile<false>(st.regs, ds_s, sm + BwdSmem::e1, sm + BwdSmem::e2, nullptr);   // E' = ds[p] e[p,:]
        {   // mask tile: row = this thread's point, 8 chunks of 8 hidden units
            uint8_t* row = sm + BwdSmem::mt + tile_off(tid, 0, kLboP);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint32_t b = (uint32_t)(bits >> (8 * c)) & 0xffu;
                uint4 v;
                v.x = ((b & 1u) ? 0x3f80u : 0u) | ((b & 2u) ? 0x3f800000u : 0u);
                v.y = ((b & 4u) ? 0x3f80u : 0u) | ((b & 8u) ? 0x3f800000u : 0u);
                v.z = ((b & 16u) ? 0x3f80u : 0u) | ((b & 32u) ? 0x3f800000u : 0u);
                v.w = ((b & 64u) ? 0x3f80u : 0u) | ((b & 128u) ? 0x3f800000u : 0u);
                *reinterpret_cast<uint4*>(row + c * kLboP) = v;
            }
        }
        fetch(st, tile + 2 * (size_t)gridDim.x);
        fence_async_smem();
        tc_fence_before();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            // dE = M (128 x 64) . W1s (64 x 32): K = 64 hidden units = 4 x 16
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const uint64_t da = umma_desc(mt + ks * 2 * kLboP, kLboP, 128);
                const uint64_t dba = umma_desc(wa + ks * 256, 128, kLboW);
                const uint64_t dbb = umma_desc(wb + ks * 256, 128, kLboW);
                umma_bf16(t_de, da, dba, idesc_de, ks > 0);
                umma_bf16(t_de, da, dbb, idesc_de, 1);
            }
            // U += M^T (64 x 128) . E' (128 x 32): K = 128 points = 8 x 16, accumulated over all tiles of this CTA
            if (grad_mlp) {
#pragma unroll
                for (int ks = 0; ks < 8; ++ks) {
                    const uint64_t da = umma_desc(mt + ks * 256, 128, kLboP);
                    const uint64_t db1 = umma_desc(e1 + ks * 256, 128, kLboP);
                    const uint64_t db2 = umma_desc(e2 + ks * 256, 128, kLboP);
                    umma_bf16(t_u, da, db1, idesc_u, (!first || ks > 0) ? 1u : 0u);
                    umma_bf16(t_u, da, db2, idesc_u, 1);
                }
            }
            umma_commit(bar);
        }
        first = false;
        mbar_wait(bar, parity);
        parity ^= 1;
        tc_fence_after();
        {   // epilogue: dE row of this thread's point -> padded staging tile
            uint32_t r[32];
            tmem_ld32(t_de + ((uint32_t)(warp * 32) << 16), r);
            tmem_ld_wait();
            float4* dst = reinterpret_cast<float4*>(sm + BwdSmem::out + tid * kOutStride);
#pragma unroll
            for (int q = 0; q < 8; ++q)
                dst[q] = make_float4(ds * __uint_as_float(r[4 * q]), ds * __uint_as_float(r[4 * q + 1]),
                                     ds * __uint_as_float(r[4 * q + 2]), ds * __uint_as_float(r[4 * q + 3]));
        }
        tc_fence_before();
        __syncthreads();
        if (level_major) {   // write-back into the planes: thread t stores the float2 of point first + t, plane by plane
            float2* planes = reinterpret_cast<float2*>(enc);
            const size_t p = tile * (size_t)kTile + tid;
            if (p < n_points) {
#pragma unroll
                for (int l = 0; l < 16; ++l)
                    __stcs(planes + (size_t)l * n_points + p, *reinterpret_cast<const float2*>(sm + BwdSmem::out + tid * kOutStride + l * 8));
            }
        } else {   // coalesced write-back (in place over the encodings of this tile)
            float4* base = reinterpret_cast<float4*>(enc) + tile * (size_t)(kTile * 8);
            const size_t first_p = tile * (size_t)kTile;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int row = i * 16 + (tid >> 3);
                const float4 v = *reinterpret_cast<const float4*>(sm + BwdSmem::out + row * kOutStride + (tid & 7) * 16);
                if (first_p + row < n_points) base[i * kThreads + tid] = v;
            }
        }
        tile += gridDim.x;
    };
    while (tile < n_tiles) {
        process(st_a);
        if (tile >= n_tiles) break;
        process(st_b);
    }
    // weight gradients from U (all MMAs of this CTA have completed: the last tile's commit was waited on)
    if (grad_mlp && !first) {
        tc_fence_after();
        uint32_t r[32];
        tmem_ld32(t_u + ((uint32_t)(warp * 32) << 16), r);
        tmem_ld_wait();
        if (lane < 16) {
            const int j = warp * 16 + lane;
            const float w2j = __ldg(w2 + j);
            float dw2 = 0.0f;
#pragma unroll
            for (int k = 0; k < LB_MLP_IN; ++k) {
                const float u = __uint_as_float(r[k]);
                atomicAdd(grad_mlp + j * LB_MLP_IN + k, w2j * u);
                dw2 = fmaf(__ldg(mlp + j * LB_MLP_IN + k), u, dw2);
            }
            atomicAdd(grad_mlp + LB_MLP_W * LB_MLP_IN + j, dw2);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, kTmemCols);
}

// ------------------------------------------------------------------------------------------
// fused backward: MLP backward + hash-grid scatter in ONE warp-specialised kernel.  dL/denc goes from tensor memory
// through a shared-memory ring straight into the scatter warps: it is never written to (or re-read from) HBM, and the
// HBM / tensor-core work of the MLP overlaps the L2-atomic-bound scatter.  One CTA of 768 threads per SM:
//   warps 0-3  (producer): mlp_bwd_umma_kernel's tile loop -- cp.async-staged raw encoding tile -> E' = ds e as bf16
//              pieces + mask tile -> MMAs (dE into TMEM, U accumulating) -> epilogue writes ds * dE rows into a ring slot
//              and arrives on the slot's "full" mbarrier
//   warps 4-27 (consumers; 4-23 with the dL/dx code): encode_bwd_kernel's per-warp work (16 points x 2 lanes, lb_encode.cuh); warp w takes the
//              16-row units w, w + 24, ... in tile order, waits for the unit's slot, copies its rows to registers,
//              releases the slot ("empty" mbarrier, 8 units per tile) and scatters.
// A symmetric version (every warp alternating between the two phases behind CTA-wide barriers) was measured slower
// than the two separate kernels (3.03 vs 2.88 ms): the scatter work per warp is data dependent and the barriers made
// every warp wait for the slowest one on every tile.
// ------------------------------------------------------------------------------------------
constexpr int kRing = 3;
constexpr int kProducerThreads = 128;
// consumer warps: 24 when the scatter fits 72 registers (table gradient only), 20 with the dL/dx code (80 registers)
template <bool kInputGrad> struct FusedShape {
    static constexpr int kConsumerWarps = kInputGrad ? 20 : 24;
    static constexpr int kThreads = kProducerThreads + 32 * kConsumerWarps;
};
struct FusedSmem {
    static constexpr uint32_t mt = 0, e1 = mt + kMaskT, e2 = e1 + kPieceP, wa = e2 + kPieceP, wb = wa + kPieceW,
                              raw = wb + kPieceW, ring = raw + 2 * kStage, bars = ring + kRing * kTile * kOutStride,
                              tmem = bars + 8 * (1 + 2 * kRing), total = tmem + 8;
};

__device__ __forceinline__ void producer_barrier() { asm volatile("bar.sync 1, 128;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// Monotonic "tiles consumed" counter in shared memory: how the gather warps of the fused forward kernels learn that a ring
// slot is free.  A count instead of an mbarrier phase because a gather warp does not take part in every use of a slot
// (28 or 20 warps, 8 units per tile, 2 slots), and a parity wait can only tell the current phase from the one before it:
// asked about phase n - 1 while the barrier is still in phase n - 3 it passes at once.  The consumers (4 MLP warps) take the
// tiles in order and add 1 each (release) when they are done with a tile's rows; tile kt may be written once
// 4 * (kt - ring + 1) increments are visible (acquire).
#ifndef LB_COUNTER_SLEEP_NS
#define LB_COUNTER_SLEEP_NS 64
#endif
__device__ __forceinline__ void counter_release_add(uint32_t addr) {
    asm volatile("red.release.cta.shared::cta.add.u32 [%0], 1;" ::"r"(addr) : "memory");
}
// called convergently by a whole warp: lane 0 polls (with a short sleep between polls, so that waiting warps do not take
// issue slots and shared-memory cycles from the working ones), __syncwarp hands the acquired view to the other lanes
__device__ __forceinline__ void counter_wait(uint32_t addr, uint32_t target) {
    if ((threadIdx.x & 31) == 0) {
        uint32_t v;
        asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
        while (v < target) {
            __nanosleep(LB_COUNTER_SLEEP_NS);
            asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
        }
    }
    __syncwarp();
}

template <bool kFromRays, bool kInputGrad, int kNDense>
__global__ void __launch_bounds__(FusedShape<kInputGrad>::kThreads, 1)
sigma_bwd_fused_kernel(lb_level_table T, const float* __restrict__ rays, const float* __restrict__ zv,
                       const float* __restrict__ x01, const float* __restrict__ mlp, const float2* __restrict__ table,
                       const float* __restrict__ enc, const float* __restrict__ d_sigma,
                       const unsigned long long* __restrict__ act, size_t n_points, int n_samples,
                       float* __restrict__ grad_mlp, float* __restrict__ grad_table, float* __restrict__ d_rays,
                       float* __restrict__ d_x01) {
    extern __shared__ __align__(128) uint8_t sm[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t bar_mma = smem_u32(sm + FusedSmem::bars);
    const uint32_t bar_full = bar_mma + 8, bar_empty = bar_full + 8 * kRing;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + FusedSmem::tmem);
    const size_t n_tiles = (n_points + kTile - 1) / kTile;

    if (warp == 0) tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    if (tid == 32) {
        mbar_init(bar_mma, 1);
        for (int i = 0; i < kRing; ++i) {
            mbar_init(bar_full + 8 * i, kProducerThreads);
            mbar_init(bar_empty + 8 * i, 8);
        }
    }
    if (tid < kProducerThreads) {                                    // W1s[j,k] = w2[j] W1[j,k] as two bf16 pieces
        const float* w2 = mlp + LB_MLP_W * LB_MLP_IN;
        for (int it = tid; it < LB_MLP_W * 4; it += kProducerThreads) {
            const int j = it >> 2, c = it & 3;
            const float sc = __ldg(w2 + j);
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = __ldg(mlp + j * LB_MLP_IN + c * 8 + e) * sc;
            uint32_t b1[4], b2[4];
            split8<false>(v, b1, b2, nullptr);
            *reinterpret_cast<uint4*>(sm + FusedSmem::wa + tile_off(j, c, kLboW)) = make_uint4(b1[0], b1[1], b1[2], b1[3]);
            *reinterpret_cast<uint4*>(sm + FusedSmem::wb + tile_off(j, c, kLboW)) = make_uint4(b2[0], b2[1], b2[2], b2[3]);
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp < 4) {
        // =============================== producer: MLP backward ===============================
        const int t = tid;                                           // row of the tile = TMEM lane
        const float* w2 = mlp + LB_MLP_W * LB_MLP_IN;
        const uint32_t t_de = tmem, t_u = tmem + 32;
        const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
        const uint32_t mt = smem_u32(sm + FusedSmem::mt), e1 = smem_u32(sm + FusedSmem::e1), e2 = smem_u32(sm + FusedSmem::e2);
        const uint32_t wa = smem_u32(sm + FusedSmem::wa), wb = smem_u32(sm + FusedSmem::wb);
        const uint32_t raw = smem_u32(sm + FusedSmem::raw);
        constexpr uint32_t idesc_de = umma_idesc(128, LB_MLP_IN, 0, 1);
        constexpr uint32_t idesc_u = umma_idesc(64, LB_MLP_IN, 1, 1);
        const uint32_t rounds = fused_rounds(n_tiles);
        if (rounds > 0) stage_tile(enc, fused_tile(0), n_points, raw);
        cp_async_commit();
        uint32_t parity = 0;
        for (uint32_t kt = 0; kt < rounds; ++kt) {
            const size_t tile = fused_tile(kt);
            const size_t p = tile * (size_t)kTile + t;
            float ds = 0.0f;
            unsigned long long bits = 0ull;
            if (p < n_points) { ds = __ldg(d_sigma + p); bits = __ldg(act + p); }
            if (kt + 1 < rounds) stage_tile(enc, fused_tile(kt + 1), n_points, raw + ((kt + 1) & 1u) * kStage);
            cp_async_commit();
            cp_async_wait<1>();
            producer_barrier();                                      // this tile's raw rows have landed
            {
                const uint8_t* row = sm + FusedSmem::raw + (kt & 1u) * kStage + t * kRowStride;
#pragma unroll
                for (int c = 0; c < 4; ++c) {                        // E' = ds e, four 8-column chunks
                    const float4 a = *reinterpret_cast<const float4*>(row + c * 32);
                    const float4 b = *reinterpret_cast<const float4*>(row + c * 32 + 16);
                    const float v[8] = {ds * a.x, ds * a.y, ds * a.z, ds * a.w, ds * b.x, ds * b.y, ds * b.z, ds * b.w};
                    uint32_t b1[4], b2[4];
                    split8<false>(v, b1, b2, nullptr);
                    const uint32_t off = tile_off(t, c, kLboP);
                    *reinterpret_cast<uint4*>(sm + FusedSmem::e1 + off) = make_uint4(b1[0], b1[1], b1[2], b1[3]);
                    *reinterpret_cast<uint4*>(sm + FusedSmem::e2 + off) = make_uint4(b2[0], b2[1], b2[2], b2[3]);
                }
#pragma unroll
                for (int c = 0; c < 8; ++c) {                        // mask tile: 8 chunks of 8 hidden units
                    const uint32_t b = (uint32_t)(bits >> (8 * c)) & 0xffu;
                    uint4 v;
                    v.x = ((b & 1u) ? 0x3f80u : 0u) | ((b & 2u) ? 0x3f800000u : 0u);
                    v.y = ((b & 4u) ? 0x3f80u : 0u) | ((b & 8u) ? 0x3f800000u : 0u);
                    v.z = ((b & 16u) ? 0x3f80u : 0u) | ((b & 32u) ? 0x3f800000u : 0u);
                    v.w = ((b & 64u) ? 0x3f80u : 0u) | ((b & 128u) ? 0x3f800000u : 0u);
                    *reinterpret_cast<uint4*>(sm + FusedSmem::mt + tile_off(t, c, kLboP)) = v;
                }
            }
            fence_async_smem();
            tc_fence_before();
            producer_barrier();
            if (tid == 0) {
                tc_fence_after();
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {                     // dE = M (128 x 64) . W1s (64 x 32)
                    const uint64_t da = umma_desc(mt + ks * 2 * kLboP, kLboP, 128);
                    umma_bf16(t_de, da, umma_desc(wa + ks * 256, 128, kLboW), idesc_de, ks > 0);
                    umma_bf16(t_de, da, umma_desc(wb + ks * 256, 128, kLboW), idesc_de, 1);
                }
                if (grad_mlp) {                                      // U += M^T (64 x 128) . E' (128 x 32)
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks) {
                        const uint64_t da = umma_desc(mt + ks * 256, 128, kLboP);
                        umma_bf16(t_u, da, umma_desc(e1 + ks * 256, 128, kLboP), idesc_u, (kt > 0 || ks > 0) ? 1u : 0u);
                        umma_bf16(t_u, da, umma_desc(e2 + ks * 256, 128, kLboP), idesc_u, 1);
                    }
                }
                umma_commit(bar_mma);
            }
            const uint32_t slot = kt % kRing;
            if (kt >= kRing) mbar_wait(bar_empty + 8 * slot, ((kt / kRing) - 1) & 1u);   // consumers released this slot
            mbar_wait(bar_mma, parity);
            parity ^= 1;
            tc_fence_after();
            {   // epilogue: ds * dE row t -> ring slot
                uint32_t r[32];
                tmem_ld32(t_de + lane_off, r);
                tmem_ld_wait();
                float4* dst = reinterpret_cast<float4*>(sm + FusedSmem::ring + slot * (kTile * kOutStride) + t * kOutStride);
#pragma unroll
                for (int qd = 0; qd < 8; ++qd)
                    dst[qd] = make_float4(ds * __uint_as_float(r[4 * qd]), ds * __uint_as_float(r[4 * qd + 1]),
                                          ds * __uint_as_float(r[4 * qd + 2]), ds * __uint_as_float(r[4 * qd + 3]));
            }
            tc_fence_before();
            mbar_arrive(bar_full + 8 * slot);                        // release: the row above is visible to the waiters
        }
        cp_async_wait<0>();
        producer_barrier();
        if (grad_mlp && blockIdx.x < n_tiles) {                      // weight gradients from U (M = 64 layout)
            tc_fence_after();
            uint32_t r[32];
            tmem_ld32(t_u + lane_off, r);
            tmem_ld_wait();
            if (lane < 16) {
                const int j = warp * 16 + lane;
                const float w2j = __ldg(w2 + j);
                float dw2 = 0.0f;
#pragma unroll
                for (int k = 0; k < LB_MLP_IN; ++k) {
                    const float u = __uint_as_float(r[k]);
                    atomicAdd(grad_mlp + j * LB_MLP_IN + k, w2j * u);
                    dw2 = fmaf(__ldg(mlp + j * LB_MLP_IN + k), u, dw2);
                }
                atomicAdd(grad_mlp + LB_MLP_W * LB_MLP_IN + j, dw2);
            }
        }
    } else {
        // =============================== consumers: hash-grid scatter ===============================
        const int cw = warp - 4;
        const size_t my_tiles = fused_rounds(n_tiles);
        for (size_t u = cw; u < my_tiles * 8; u += FusedShape<kInputGrad>::kConsumerWarps) {
            const uint32_t kt = (uint32_t)(u >> 3);
            const int row = (int)(u & 7) * 16 + (lane >> 1);
            const size_t tile = fused_tile(kt);
            const size_t p = tile * (size_t)kTile + row;
            const bool valid = p < n_points;
            lb_enc::Point q = {0.5f, 0.5f, 0.5f};
            float zval = 0.0f;
            if (valid) q = lb_enc::load_point<kFromRays>(rays, zv, x01, p, n_samples, &zval);
            const uint32_t slot = kt % kRing;
            mbar_wait(bar_full + 8 * slot, (kt / kRing) & 1u);
            float de[LB_MLP_IN];
            {
                const float4* src = reinterpret_cast<const float4*>(sm + FusedSmem::ring + slot * (kTile * kOutStride) + row * kOutStride);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 v = src[i];
                    de[4 * i] = v.x; de[4 * i + 1] = v.y; de[4 * i + 2] = v.z; de[4 * i + 3] = v.w;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_empty + 8 * slot);        // this unit's rows are in registers
            lb_enc::encode_bwd_point<kFromRays, kInputGrad, kNDense, lb_enc::kAggLevelsFused, lb_enc::kAggMaxRunsFused>(
                T, q, zval, de, valid, p, n_samples, table, grad_table, d_rays, d_x01, 0, T.n_levels);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, kTmemCols);
}

template <bool kFromRays, bool kInputGrad, int kNDense>
int launch_fused(const lb_plan* plan, const lb_level_table& T, const float* rays, const float* z, const float* x01,
                 const float* params, const float* enc, const uint64_t* act, const float* d_sigma, size_t n_points,
                 int n_samples, float* grad_params, float* d_rays, float* d_x01, cudaStream_t st) {
    auto kern = sigma_bwd_fused_kernel<kFromRays, kInputGrad, kNDense>;
    int rc = lb_check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FusedSmem::total),
                           "sigma_bwd_fused smem opt-in");
    if (rc) return rc;
    const size_t tiles = (n_points + kTile - 1) / kTile;
    const size_t cap = (size_t)plan->sm_count;
    kern<<<(int)(tiles < cap ? tiles : cap), FusedShape<kInputGrad>::kThreads, FusedSmem::total, st>>>(
        T, rays, z, x01, params, reinterpret_cast<const float2*>(params + LB_MLP_PARAMS), enc, d_sigma,
        reinterpret_cast<const unsigned long long*>(act), n_points, n_samples, grad_params,
        grad_params ? grad_params + LB_MLP_PARAMS : nullptr, d_rays, d_x01);
    return lb_check_cuda(cudaGetLastError(), "sigma_bwd_fused launch");
}

// ------------------------------------------------------------------------------------------
// fused forward: hash-grid gather + MLP in one launch.
//
// 28 gather warps (the 2-lanes-per-point walk of lb_encode.cu, 16 points per warp-unit) write their rows into a
// 2-tile shared-memory ring (and to `enc` for the backward when it is non-NULL); 4 MLP warps take each full tile
// through the tf32 / TMEM-operand pipeline of mlp_fwd_umma_kernel.  The hand-off never leaves the SM, so the MLP's
// re-read of the encodings (128 B per point) and its launch disappear behind the gather, which stays bound by L2
// sector fills.  A / H are double-buffered in TMEM: the MMAs of tile k run under the epilogue of tile k - 1.
// ------------------------------------------------------------------------------------------
// Ring depth: the gather warps issue their loads before they wait for a slot, so depth buys nothing (2 / 3 / 4 / 6 / 8
// slots: lb_sigma_fwd 3.16 / 3.19 / 3.19 / 3.27 / 3.36 ms per bench step) while every 18 KB slot comes out of the L1 that
// serves the gathers; two slots fit the 64 KB shared-memory carve-out.
#ifndef LB_FWD_RING
#define LB_FWD_RING 2
#endif
constexpr int kFwdRing = LB_FWD_RING;
// Gather warps of the fused forward: 28 + the 4 MLP warps = 1024 threads, the CTA maximum, at 64 registers (same spills
// as 24 warps at 72).  The gathers are latency-bound requests into an L1TEX that is 84 % busy: more warps in flight,
// measured A-B-A-B inside the bench step: lb_sigma_fwd 3.159 -> 3.108 ms, 0.908 -> 0.922 of the L2 gather ceiling.
#ifndef LB_GATHER_WARPS
#define LB_GATHER_WARPS 28
#endif
constexpr int kGatherWarps = LB_GATHER_WARPS;
constexpr int kFusedFwdThreads = kProducerThreads + 32 * kGatherWarps;
constexpr uint32_t kFusedFwdTmemCols = 256;        // H0 64 | H1 64 | A0 hi 32 lo 32 | A1 hi 32 lo 32
struct FusedFwdSmem {
    static constexpr uint32_t ring = 0, whi = ring + kFwdRing * kStage, wlo = whi + 8 * kLboW, w2 = wlo + 8 * kLboW,
                              v = w2 + 64 * 4, bars = v + 32 * 4, tmem = bars + 8 * (2 + 2 * kFwdRing), total = tmem + 8;
};

template <bool kFromRays, int kNDense>
__global__ void __launch_bounds__(kFusedFwdThreads, 1)
sigma_fwd_fused_kernel(lb_level_table T, const float* __restrict__ rays, const float* __restrict__ zv,
                       const float* __restrict__ x01, const float* __restrict__ mlp, const float2* __restrict__ table,
                       size_t n_points, int n_samples, float* __restrict__ sigma, float* __restrict__ enc,
                       unsigned long long* __restrict__ act) {
    extern __shared__ __align__(128) uint8_t sm[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t bar_mma = smem_u32(sm + FusedFwdSmem::bars);
    const uint32_t bar_full = bar_mma + 16, done_cnt = bar_full + 8 * kFwdRing;    // counter_wait / counter_release_add
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + FusedFwdSmem::tmem);
    float* w2s = reinterpret_cast<float*>(sm + FusedFwdSmem::w2);
    float* vs = reinterpret_cast<float*>(sm + FusedFwdSmem::v);
    const size_t n_tiles = (n_points + kTile - 1) / kTile;
    const size_t my_tiles = fused_rounds(n_tiles);

    if (warp == 0) tmem_alloc(smem_u32(tmem_slot), kFusedFwdTmemCols);
    if (tid == 32) {
        mbar_init(bar_mma, 1);
        mbar_init(bar_mma + 8, 1);
        for (int i = 0; i < kFwdRing; ++i) mbar_init(bar_full + 8 * i, 8);
        *reinterpret_cast<uint32_t*>(sm + FusedFwdSmem::bars + 16 + 8 * kFwdRing) = 0u;     // done_cnt
    }
    if (tid < kProducerThreads) {
        convert_weights_tf32(mlp, sm + FusedFwdSmem::whi, sm + FusedFwdSmem::wlo);
        if (tid < LB_MLP_W) w2s[tid] = 0.5f * __ldg(mlp + LB_MLP_W * LB_MLP_IN + tid);
        if (tid >= 64 && tid < 64 + LB_MLP_IN) {                     // v = W1^T w2 / 2 (see mlp_fwd_umma_kernel)
            const int k = tid - 64;
            float a = 0.0f;
            for (int j = 0; j < LB_MLP_W; ++j)
                a = fmaf(0.5f * __ldg(mlp + LB_MLP_W * LB_MLP_IN + j), __ldg(mlp + j * LB_MLP_IN + k), a);
            vs[k] = a;
        }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp < 4) {
        // =============================== MLP warps ===============================
        const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
        const uint32_t bhi = smem_u32(sm + FusedFwdSmem::whi), blo = smem_u32(sm + FusedFwdSmem::wlo);
        constexpr uint32_t idesc = umma_idesc_tf32(128, LB_MLP_W);
        float lin_prev = 0.0f;
        for (uint32_t kt = 0; kt <= my_tiles; ++kt) {
            const uint32_t b = kt & 1u;
            float lin = 0.0f;
            if (kt < my_tiles) {
                const uint32_t slot = kt % kFwdRing;
                const uint32_t t_ahi = tmem + 128 + b * 64, t_alo = t_ahi + 32;
                mbar_wait(bar_full + 8 * slot, (kt / kFwdRing) & 1u);
                const uint8_t* row = sm + FusedFwdSmem::ring + slot * kStage + tid * kRowStride;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    uint32_t hi[16], lo[16];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 v = *reinterpret_cast<const float4*>(row + (h * 4 + q) * 16);
                        const float4 c = *reinterpret_cast<const float4*>(vs + (h * 4 + q) * 4);
                        lin = fmaf(v.x, c.x, lin); lin = fmaf(v.y, c.y, lin);
                        lin = fmaf(v.z, c.z, lin); lin = fmaf(v.w, c.w, lin);
                        uint4 a, d;
                        split_tf32(v, a, d);
                        hi[4 * q] = a.x; hi[4 * q + 1] = a.y; hi[4 * q + 2] = a.z; hi[4 * q + 3] = a.w;
                        lo[4 * q] = d.x; lo[4 * q + 1] = d.y; lo[4 * q + 2] = d.z; lo[4 * q + 3] = d.w;
                    }
                    tmem_st16(t_ahi + lane_off + h * 16, hi);
                    tmem_st16(t_alo + lane_off + h * 16, lo);
                }
                __syncwarp();
                if (lane == 0) counter_release_add(done_cnt);        // the rows of this warp are in registers / TMEM
                tmem_st_wait();
            }
            tc_fence_before();
            producer_barrier();        // A[b] complete; every thread's epilogue of tile kt - 2 (the last user of H[b]) is done
            if (tid == 0 && kt < my_tiles) {
                tc_fence_after();
                const uint32_t t_h = tmem + b * 64, t_ahi = tmem + 128 + b * 64, t_alo = t_ahi + 32;
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const uint64_t dbh = umma_desc(bhi + ks * 2 * kLboW, kLboW, 128);
                    const uint64_t dbl = umma_desc(blo + ks * 2 * kLboW, kLboW, 128);
                    umma_tf32_ts(t_h, t_ahi + ks * 8, dbh, idesc, ks > 0);
                    umma_tf32_ts(t_h, t_ahi + ks * 8, dbl, idesc, 1);
                    umma_tf32_ts(t_h, t_alo + ks * 8, dbh, idesc, 1);
                }
                umma_commit(bar_mma + 8 * b);
            }
            if (kt > 0) {              // epilogue of tile kt - 1 while the tensor pipe works on tile kt
                const uint32_t pb = b ^ 1u, pk = kt - 1;
                mbar_wait(bar_mma + 8 * pb, (pk >> 1) & 1u);
                tc_fence_after();
                float s = lin_prev;
                unsigned long long bits = 0ull;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                    uint32_t r[32];
                    tmem_ld32(tmem + pb * 64 + lane_off + half * 32, r);
                    tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 2; ++g) {
                        float mf = 0.0f;
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const float h = __uint_as_float(r[g * 16 + j]);
                            s = fmaf(w2s[half * 32 + g * 16 + j], fabsf(h), s);
                            mf = fmaf(__saturatef(h * 8.507059173e37f), (float)(1u << j), mf);
                        }
                        const uint32_t m16 = __float_as_uint(mf + 8388608.0f) & 0xffffu;
                        bits |= (unsigned long long)m16 << (32 * half + 16 * g);
                    }
                }
                const size_t p = fused_tile(pk) * (size_t)kTile + tid;
                if (p < n_points) {
                    sigma[p] = s;
                    if (act) act[p] = bits;
                }
            }
            lin_prev = lin;
        }
    } else {
        // =============================== gather warps ===============================
        const int cw = warp - 4, half = lane & 1;
        for (size_t u = cw; u < my_tiles * 8; u += kGatherWarps) {
            const uint32_t kt = (uint32_t)(u >> 3);
            const int row = (int)(u & 7) * 16 + (lane >> 1);
            const size_t p = fused_tile(kt) * (size_t)kTile + row;
            const bool valid = p < n_points;
            lb_enc::Point q = {0.5f, 0.5f, 0.5f};
            if (valid) q = lb_enc::load_point<kFromRays>(rays, zv, x01, p, n_samples, nullptr);
            float out[16];
            lb_enc::encode_fwd_point<kNDense>(T, q, half, table, 0, 16, out);
            const uint32_t slot = kt % kFwdRing;
            if (kt >= kFwdRing) counter_wait(done_cnt, 4u * (kt - kFwdRing + 1));      // tile kt - ring has been consumed
            float4* dst = reinterpret_cast<float4*>(sm + FusedFwdSmem::ring + slot * kStage + row * kRowStride + half * 64);
#pragma unroll
            for (int i = 0; i < 4; ++i) dst[i] = make_float4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]);
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_full + 8 * slot);
            if (enc && valid) {
                float4* g = reinterpret_cast<float4*>(enc + p * LB_MLP_IN + half * 16);
#pragma unroll
                for (int i = 0; i < 4; ++i) __stcs(g + i, make_float4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, kFusedFwdTmemCols);
}

template <bool kFromRays, int kNDense>
int launch_fused_fwd_src(const lb_plan* plan, const lb_level_table& T, const float* rays, const float* z, const float* x01,
                         const float* params, size_t n_points, int n_samples, float* sigma, float* enc, uint64_t* act,
                         cudaStream_t st) {
    auto kern = sigma_fwd_fused_kernel<kFromRays, kNDense>;
    int rc = lb_check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FusedFwdSmem::total),
                           "sigma_fwd_fused smem opt-in");
    if (rc) return rc;
    const size_t tiles = (n_points + kTile - 1) / kTile;
    const size_t cap = (size_t)plan->sm_count;
    kern<<<(int)(tiles < cap ? tiles : cap), kFusedFwdThreads, FusedFwdSmem::total, st>>>(
        T, rays, z, x01, params, reinterpret_cast<const float2*>(params + LB_MLP_PARAMS), n_points, n_samples, sigma, enc,
        reinterpret_cast<unsigned long long*>(act));
    return lb_check_cuda(cudaGetLastError(), "sigma_fwd_fused launch");
}

// point source: (rays, z) -> x = o + d z inside the kernel, or x01 [P,3] as handed to tcnn.NetworkWithInputEncoding
template <int kNDense>
int launch_fused_fwd(const lb_plan* plan, const lb_level_table& T, const float* rays, const float* z, const float* x01,
                     const float* params, size_t n_points, int n_samples, float* sigma, float* enc, uint64_t* act,
                     cudaStream_t st) {
    return rays ? launch_fused_fwd_src<true, kNDense>(plan, T, rays, z, nullptr, params, n_points, n_samples, sigma, enc, act, st)
                : launch_fused_fwd_src<false, kNDense>(plan, T, nullptr, nullptr, x01, params, n_points, 1, sigma, enc, act, st);
}

// ------------------------------------------------------------------------------------------
// fused forward with the forward-mode input derivative: sigma AND d sigma / d x01 [P,3] in one launch.
//
// Pose refinement needs dL/d(ray) = sum over the ray's samples of dL/dsigma * d sigma/d x * dx/d(ray).  The reverse-mode
// route (sigma_bwd_fused_kernel<.., kInputGrad = true>) gathers the 8 corners of all 16 levels a second time in the
// backward: 42 more L2 sectors per point in a kernel that is already bound by its 54 reduction sectors (8.1 instead of
// 5.1 ms per step).  Forward mode needs no second gather: the gather warps already hold the corner values, so they also
// form the encoding's Jacobian J = d e / d x01 (32 x 3 per point, lb_encode.cuh); once the MLP warps know the point's ReLU
// mask m they run the backward's first MMA, dE = M . (w2 (.) W1) (mask tile in shared memory, bf16 two-piece weights, TMEM
// accumulator -- the same arithmetic as sigma_bwd_fused_kernel), and contract: d sigma / d x01 = J^T dE.  12 bytes per
// point leave the SM instead of a 384-byte Jacobian, and the backward is the table-gradient-only kernel plus a per-ray
// reduction (lb_field.cu).
//
// Where J waits for dE: TENSOR MEMORY.  H 64 + A 64 + dE 32 columns are in use, a tile's Jacobian is 128 lanes x 96
// columns, so three tiles fit the remaining 352 columns.  A warp may only touch the TMEM lanes of its own quarter (warp
// id % 4) and a thread only its own lane, so the Jacobian is stored the way the gather holds it -- lane pair (2k, 2k + 1)
// of a 16-point unit keeps the feature-0 / feature-1 triples of point k -- and the two units of a 32-row quarter share its
// 32 lanes side by side: unit u = 2q + h of a tile writes columns [48 h, 48 h + 48) of lanes 32 q + (0..31), 12 columns
// per four levels straight from registers (tcgen05.st).  The MLP thread of row t then finds in ITS lane the triples of
// rows 32 w + 16 h + (t >> 1) (feature t & 1): it fetches those rows' dE by warp shuffle, forms the partial dot products,
// adds the two features (lane pair) and sends the sum to the row's own thread.  24 gather warps: warp g serves quarter
// g & 3, unit half (g >> 2) & 1, ring slot g >> 3 (tiles slot, slot + 3, ...), so all three tiles in flight are always
// being filled.  Only the encoding rows pass through shared memory (3 x 18 KB; 97 KB in all).
//
// Pipeline per 128-point tile (MLP warps, thread t = point t): E row -> tf32 hi / lo -> TMEM; MMA H; epilogue (sigma, mask
// bits) -> mask tile; MMA dE; TMEM -> registers; the contraction above; release the slot (monotonic counter, see
// counter_wait).  No double buffering of the two MMA round trips.
//
// Measured inside the c3 bench step (lb_sigma_fwd_jvp per step, two 16384 x 512 chunks; the plain forward: 3.16 ms), every
// variant passing the same parity tests (profiles/README.md section 4):
//   J rows in shared memory next to the E rows (2 slots x 68 KB = 256 points in flight), 12 / 16 / 20 / 24 gather warps:
//       4.56 / 4.37 / 4.57 / 4.75 ms (waiting warps spinning instead of sleeping: 16 / 20 / 24 = 4.48 / 4.96 / 5.29)
//   J in tensor memory, ONE lane per point (all 8 corners per lane, no exchanges, 12 warps x 32 points, 128 registers):
//       5.43 ms -- 3 gather warps per scheduler cannot hide the latency of the 8-corner lane
//   J in tensor memory, lane pair per point as above (24 warps, 72 registers):                      4.27 ms  <- this kernel
//   the same with the pair splitting the work by FEATURE (4-byte loads of all 8 corners, no shuffles at all): 4.47 ms
// ncu (profiles/r02_ncu_jvp_fwd.csv): the same 50.3 tex sectors per point as the plain forward, L2 44 % busy; the LSU data
// pipe is at 79 % of its peak (global-load wavefronts + the 3 extra shuffles per level), issue 59 %, and the gather warps
// still spend time asleep waiting for a slot: the serial MLP chain of a tile (two MMA round trips + 64 shuffles) holds it.
// ------------------------------------------------------------------------------------------
constexpr int kJvpGatherWarps = 24;
constexpr int kJvpThreads = kProducerThreads + 32 * kJvpGatherWarps;
constexpr int kJvpRing = 3;
constexpr uint32_t kJvpTmemCols = 512;              // H 64 | A hi 32 | A lo 32 | dE 32 | J0 96 | J1 96 | J2 96 (448 used)
constexpr uint32_t kJvpColJ = 160;
struct JvpSmem {
    static constexpr uint32_t ring = 0, mt = ring + kJvpRing * kStage, wa = mt + kMaskT, wb = wa + kPieceW,
                              whi = wb + kPieceW, wlo = whi + 8 * kLboW, w2 = wlo + 8 * kLboW, v = w2 + 64 * 4,
                              bars = v + 32 * 4, done = bars + 8 * (2 + kJvpRing), tmem = done + 8, total = tmem + 8;
};

template <bool kFromRays, int kNDense>
__global__ void __launch_bounds__(kJvpThreads, 1)
sigma_fwd_jvp_fused_kernel(lb_level_table T, const float* __restrict__ rays, const float* __restrict__ zv,
                          const float* __restrict__ x01, const float* __restrict__ mlp, const float2* __restrict__ table,
                          size_t n_points, int n_samples, float* __restrict__ sigma, float* __restrict__ enc,
                          unsigned long long* __restrict__ act, float* __restrict__ dsigma_dx) {
    extern __shared__ __align__(128) uint8_t sm[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t bar_h = smem_u32(sm + JvpSmem::bars), bar_de = bar_h + 8;
    const uint32_t bar_full = bar_h + 16, done_cnt = smem_u32(sm + JvpSmem::done);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + JvpSmem::tmem);
    float* w2s = reinterpret_cast<float*>(sm + JvpSmem::w2);
    float* vs = reinterpret_cast<float*>(sm + JvpSmem::v);
    const size_t n_tiles = (n_points + kTile - 1) / kTile;
    const size_t my_tiles = fused_rounds(n_tiles);

    if (warp == 0) tmem_alloc(smem_u32(tmem_slot), kJvpTmemCols);
    if (tid == 32) {
        mbar_init(bar_h, 1);
        mbar_init(bar_de, 1);
        for (int i = 0; i < kJvpRing; ++i) mbar_init(bar_full + 8 * i, 8);      // eight 16-point units per tile
        *reinterpret_cast<uint32_t*>(sm + JvpSmem::done) = 0u;
    }
    if (tid < kProducerThreads) {
        convert_weights_tf32(mlp, sm + JvpSmem::whi, sm + JvpSmem::wlo);
        const float* w2 = mlp + LB_MLP_W * LB_MLP_IN;
        if (tid < LB_MLP_W) w2s[tid] = 0.5f * __ldg(w2 + tid);
        if (tid >= 64 && tid < 64 + LB_MLP_IN) {
            const int k = tid - 64;
            float a = 0.0f;
            for (int j = 0; j < LB_MLP_W; ++j) a = fmaf(0.5f * __ldg(w2 + j), __ldg(mlp + j * LB_MLP_IN + k), a);
            vs[k] = a;
        }
        for (int it = tid; it < LB_MLP_W * 4; it += kProducerThreads) {
            const int j = it >> 2, c = it & 3;
            const float sc = __ldg(w2 + j);
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = __ldg(mlp + j * LB_MLP_IN + c * 8 + e) * sc;
            uint32_t b1[4], b2[4];
            split8<false>(v, b1, b2, nullptr);
            *reinterpret_cast<uint4*>(sm + JvpSmem::wa + tile_off(j, c, kLboW)) = make_uint4(b1[0], b1[1], b1[2], b1[3]);
            *reinterpret_cast<uint4*>(sm + JvpSmem::wb + tile_off(j, c, kLboW)) = make_uint4(b2[0], b2[1], b2[2], b2[3]);
        }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp < 4) {
        // =============================== MLP warps ===============================
        const uint32_t lane_off = (uint32_t)(warp * 32) << 16;
        const uint32_t bhi = smem_u32(sm + JvpSmem::whi), blo = smem_u32(sm + JvpSmem::wlo);
        const uint32_t mt = smem_u32(sm + JvpSmem::mt), wa = smem_u32(sm + JvpSmem::wa), wb = smem_u32(sm + JvpSmem::wb);
        const uint32_t t_h = tmem, t_ahi = tmem + 64, t_alo = tmem + 96, t_de = tmem + 128;
        constexpr uint32_t idesc_h = umma_idesc_tf32(128, LB_MLP_W);
        constexpr uint32_t idesc_de = umma_idesc(128, LB_MLP_IN, 0, 1);
        for (uint32_t kt = 0; kt < my_tiles; ++kt) {
            const uint32_t slot = kt % kJvpRing, par = kt & 1u;
            mbar_wait(bar_full + 8 * slot, (kt / kJvpRing) & 1u);
            tc_fence_after();                                        // the gather warps' tcgen05.st of this tile's J
            const uint8_t* row = sm + JvpSmem::ring + slot * kStage + tid * kRowStride;
            float lin = 0.0f;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                uint32_t hi[16], lo[16];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 v = *reinterpret_cast<const float4*>(row + (h * 4 + q) * 16);
                    const float4 c = *reinterpret_cast<const float4*>(vs + (h * 4 + q) * 4);
                    lin = fmaf(v.x, c.x, lin); lin = fmaf(v.y, c.y, lin);
                    lin = fmaf(v.z, c.z, lin); lin = fmaf(v.w, c.w, lin);
                    uint4 a, d;
                    split_tf32(v, a, d);
                    hi[4 * q] = a.x; hi[4 * q + 1] = a.y; hi[4 * q + 2] = a.z; hi[4 * q + 3] = a.w;
                    lo[4 * q] = d.x; lo[4 * q + 1] = d.y; lo[4 * q + 2] = d.z; lo[4 * q + 3] = d.w;
                }
                tmem_st16(t_ahi + lane_off + h * 16, hi);
                tmem_st16(t_alo + lane_off + h * 16, lo);
            }
            tmem_st_wait();
            tc_fence_before();
            producer_barrier();
            if (tid == 0) {
                tc_fence_after();
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const uint64_t dbh = umma_desc(bhi + ks * 2 * kLboW, kLboW, 128);
                    const uint64_t dbl = umma_desc(blo + ks * 2 * kLboW, kLboW, 128);
                    umma_tf32_ts(t_h, t_ahi + ks * 8, dbh, idesc_h, ks > 0);
                    umma_tf32_ts(t_h, t_ahi + ks * 8, dbl, idesc_h, 1);
                    umma_tf32_ts(t_h, t_alo + ks * 8, dbh, idesc_h, 1);
                }
                umma_commit(bar_h);
            }
            mbar_wait(bar_h, par);
            tc_fence_after();
            float s = lin;
            unsigned long long bits = 0ull;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                uint32_t r[32];
                tmem_ld32(t_h + lane_off + half * 32, r);
                tmem_ld_wait();
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    float mf = 0.0f;
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float h = __uint_as_float(r[g * 16 + j]);
                        s = fmaf(w2s[half * 32 + g * 16 + j], fabsf(h), s);
                        mf = fmaf(__saturatef(h * 8.507059173e37f), (float)(1u << j), mf);
                    }
                    const uint32_t m16 = __float_as_uint(mf + 8388608.0f) & 0xffffu;
                    bits |= (unsigned long long)m16 << (32 * half + 16 * g);
                }
            }
            const size_t p = fused_tile(kt) * (size_t)kTile + tid;
            if (p < n_points) {
                sigma[p] = s;
                if (act) act[p] = bits;
            }
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const uint32_t b = (uint32_t)(bits >> (8 * c)) & 0xffu;
                uint4 v;
                v.x = ((b & 1u) ? 0x3f80u : 0u) | ((b & 2u) ? 0x3f800000u : 0u);
                v.y = ((b & 4u) ? 0x3f80u : 0u) | ((b & 8u) ? 0x3f800000u : 0u);
                v.z = ((b & 16u) ? 0x3f80u : 0u) | ((b & 32u) ? 0x3f800000u : 0u);
                v.w = ((b & 64u) ? 0x3f80u : 0u) | ((b & 128u) ? 0x3f800000u : 0u);
                *reinterpret_cast<uint4*>(sm + JvpSmem::mt + tile_off(tid, c, kLboP)) = v;
            }
            fence_async_smem();
            tc_fence_before();
            producer_barrier();
            if (tid == 0) {
                tc_fence_after();
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const uint64_t da = umma_desc(mt + ks * 2 * kLboP, kLboP, 128);
                    umma_bf16(t_de, da, umma_desc(wa + ks * 256, 128, kLboW), idesc_de, ks > 0);
                    umma_bf16(t_de, da, umma_desc(wb + ks * 256, 128, kLboW), idesc_de, 1);
                }
                umma_commit(bar_de);
            }
            mbar_wait(bar_de, par);
            tc_fence_after();
            {
                uint32_t de[32];
                tmem_ld32(t_de + lane_off, de);
                tmem_ld_wait();
                // This lane's 96 Jacobian columns: [48 h + 3 l + d] = d e[2l + f] / d x01[d] of row 32 w + 16 h + (lane >> 1),
                // f = lane & 1.  Partial dot products with that row's dE (shuffled from its thread), then f-pair sum.
                const uint32_t t_j = tmem + kJvpColJ + slot * 96 + lane_off;
                const int f = lane & 1;
                float gp[2][3] = {{0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}};
                float dv = 0.0f;
#pragma unroll
                for (int c3 = 0; c3 < 3; ++c3) {
                    uint32_t jr[32];
                    tmem_ld32(t_j + c3 * 32, jr);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const int col = c3 * 32 + i, h = col / 48, l = (col % 48) / 3, d = col % 3;
                        if (d == 0) {                                 // dE[2l + f] of row 16 h + (lane >> 1) of this quarter
                            const float d0 = __shfl_sync(0xffffffffu, __uint_as_float(de[2 * l]), 16 * h + (lane >> 1));
                            const float d1 = __shfl_sync(0xffffffffu, __uint_as_float(de[2 * l + 1]), 16 * h + (lane >> 1));
                            dv = f ? d1 : d0;
                        }
                        gp[h][d] = fmaf(dv, __uint_as_float(jr[i]), gp[h][d]);
                    }
                }
                float g[3];
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    const float s0 = gp[0][d] + __shfl_xor_sync(0xffffffffu, gp[0][d], 1);
                    const float s1 = gp[1][d] + __shfl_xor_sync(0xffffffffu, gp[1][d], 1);
                    const float r0 = __shfl_sync(0xffffffffu, s0, 2 * (lane & 15));      // row `lane` = half lane >> 4, point lane & 15
                    const float r1 = __shfl_sync(0xffffffffu, s1, 2 * (lane & 15));
                    g[d] = (lane >> 4) ? r1 : r0;
                }
                if (p < n_points) {
                    dsigma_dx[3 * p + 0] = g[0];
                    dsigma_dx[3 * p + 1] = g[1];
                    dsigma_dx[3 * p + 2] = g[2];
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) counter_release_add(done_cnt);            // this warp is done with the slot's E rows and J columns
        }
    } else {
        // =============================== gather warps ===============================
        const int g = warp - 4, q4 = g & 3, uh = (g >> 2) & 1, half = lane & 1;      // (g + 4) % 4 == q4: this warp's TMEM quarter
        const uint32_t slot = (uint32_t)g >> 3;                      // a warp always fills the same slot
        const uint32_t t_j = tmem + kJvpColJ + slot * 96 + uh * 48 + ((uint32_t)(q4 * 32) << 16);
        const int row = (2 * q4 + uh) * 16 + (lane >> 1);
        uint8_t* base = sm + JvpSmem::ring + slot * kStage;
        for (uint32_t kt = slot; kt < my_tiles; kt += kJvpRing) {
            const size_t p = fused_tile(kt) * (size_t)kTile + row;
            const bool valid = p < n_points;
            lb_enc::Point pt = {0.5f, 0.5f, 0.5f};
            if (valid) pt = lb_enc::load_point<kFromRays>(rays, zv, x01, p, n_samples, nullptr);
            float out[16];
            lb_enc::encode_fwd_jvp_point<kNDense>(T, pt, half, table, out, [&](int gi, const float* jl) {
                if (gi == 0 && kt >= kJvpRing) {                      // tile kt - 3 (this slot: J columns and E rows) is done
                    counter_wait(done_cnt, 4u * (kt - kJvpRing + 1));
                    tc_fence_after();
                }
#pragma unroll
                for (int i = 0; i < 3; ++i)
                    tmem_st4(t_j + gi * 12 + i * 4, __float_as_uint(jl[4 * i]), __float_as_uint(jl[4 * i + 1]),
                             __float_as_uint(jl[4 * i + 2]), __float_as_uint(jl[4 * i + 3]));
            });
            float4* dst = reinterpret_cast<float4*>(base + row * kRowStride + half * 64);
#pragma unroll
            for (int i = 0; i < 4; ++i) dst[i] = make_float4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_full + 8 * slot);
            if (enc && valid) {
                float4* ge = reinterpret_cast<float4*>(enc + p * LB_MLP_IN + half * 16);
#pragma unroll
                for (int i = 0; i < 4; ++i) __stcs(ge + i, make_float4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]));
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, kJvpTmemCols);
}

template <bool kFromRays, int kNDense>
int launch_fused_fwd_jvp_src(const lb_plan* plan, const lb_level_table& T, const float* rays, const float* z, const float* x01,
                             const float* params, size_t n_points, int n_samples, float* sigma, float* enc, uint64_t* act,
                             float* dsigma_dx, cudaStream_t st) {
    auto kern = sigma_fwd_jvp_fused_kernel<kFromRays, kNDense>;
    constexpr int kSmemBytes = (int)JvpSmem::total, kBlock = kJvpThreads;
    int rc = lb_check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes),
                           "sigma_fwd_jvp_fused smem opt-in");
    if (rc) return rc;
    const size_t tiles = (n_points + kTile - 1) / kTile;
    const size_t cap = (size_t)plan->sm_count;
    kern<<<(int)(tiles < cap ? tiles : cap), kBlock, kSmemBytes, st>>>(
        T, rays, z, x01, params, reinterpret_cast<const float2*>(params + LB_MLP_PARAMS), n_points, n_samples, sigma, enc,
        reinterpret_cast<unsigned long long*>(act), dsigma_dx);
    return lb_check_cuda(cudaGetLastError(), "sigma_fwd_jvp_fused launch");
}

template <int kNDense>
int launch_fused_fwd_jvp(const lb_plan* plan, const lb_level_table& T, const float* rays, const float* z, const float* x01,
                         const float* params, size_t n_points, int n_samples, float* sigma, float* enc, uint64_t* act,
                         float* dsigma_dx, cudaStream_t st) {
    return rays ? launch_fused_fwd_jvp_src<true, kNDense>(plan, T, rays, z, nullptr, params, n_points, n_samples, sigma, enc, act, dsigma_dx, st)
                : launch_fused_fwd_jvp_src<false, kNDense>(plan, T, nullptr, nullptr, x01, params, n_points, 1, sigma, enc, act, dsigma_dx, st);
}

int umma_grid(int sm_count, size_t n_points, int ctas_per_sm) {
    const size_t tiles = (n_points + kTile - 1) / kTile;
    const size_t cap = (size_t)sm_count * ctas_per_sm;
    return (int)(tiles < cap ? (tiles ? tiles : 1) : cap);
}

}  // namespace

int lb_launch_mlp_fwd_umma(const lb_plan* plan, const float* mlp, const float* enc, size_t n_points, float* sigma,
                           uint64_t* act, cudaStream_t st, bool level_major) {
    // per launch: the attribute is per device, and a process may drive several devices
    int rc = lb_check_cuda(cudaFuncSetAttribute(mlp_fwd_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                (int)FwdSmem::total), "mlp_fwd_umma smem opt-in");
    if (rc) return rc;
    mlp_fwd_umma_kernel<<<umma_grid(plan->sm_count, n_points, kFwdCtasPerSm), kThreads, FwdSmem::total, st>>>(
        mlp, enc, n_points, sigma, reinterpret_cast<unsigned long long*>(act), level_major ? 1 : 0);
    return lb_check_cuda(cudaGetLastError(), "mlp_fwd_umma launch");
}

int lb_launch_mlp_bwd_umma(const lb_plan* plan, const float* mlp, float* enc_inout, const float* d_sigma,
                           const uint64_t* act, size_t n_points, float* grad_mlp, cudaStream_t st, bool level_major) {
    int rc = lb_check_cuda(cudaFuncSetAttribute(mlp_bwd_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                (int)BwdSmem::total), "mlp_bwd_umma smem opt-in");
    if (rc) return rc;
    mlp_bwd_umma_kernel<<<umma_grid(plan->sm_count, n_points, kBwdCtasPerSm), kThreads, BwdSmem::total, st>>>(
        mlp, enc_inout, d_sigma, reinterpret_cast<const unsigned long long*>(act), n_points, grad_mlp, level_major ? 1 : 0);
    return lb_check_cuda(cudaGetLastError(), "mlp_bwd_umma launch");
}

int lb_launch_sigma_bwd_fused(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                              const float* params, const float* enc, const uint64_t* act, const float* d_sigma,
                              size_t n_points, int n_samples, float* grad_params, float* d_rays, float* d_x01,
                              cudaStream_t st) {
#ifdef LB_TEST_SWITCHES       // test builds: LONER_B200_FUSED_BWD=0 runs the two-kernel path on configurations the fused kernel covers
    static int disabled = -1;
    if (disabled < 0) {
        const char* e = getenv("LONER_B200_FUSED_BWD");
        disabled = (e && e[0] == '0') ? 1 : 0;
    }
#else
    constexpr int disabled = 0;
#endif
    const int nd = lb_encode_dense_prefix(plan);
    if (disabled || (!rays && !x01) || !act || nd < 1 || !lb_encode_single_group(plan) || plan->cfg.n_levels != 16) return 1;
    const lb_level_table T = lb_make_level_table(plan);
    const bool in_grad = d_rays || d_x01;
#define LB_FUSED_ARGS plan, T, rays, z, x01, params, enc, act, d_sigma, n_points, rays ? n_samples : 1, grad_params, d_rays, d_x01, st
#define LB_FUSED_CASE(ND)                                                                                          \
    case ND:                                                                                                       \
        if (rays) return in_grad ? launch_fused<true, true, ND>(LB_FUSED_ARGS) : launch_fused<true, false, ND>(LB_FUSED_ARGS); \
        return in_grad ? launch_fused<false, true, ND>(LB_FUSED_ARGS) : launch_fused<false, false, ND>(LB_FUSED_ARGS);
    switch (nd) {
        LB_FUSED_CASE(1)
        LB_FUSED_CASE(2)
        LB_FUSED_CASE(3)
        LB_FUSED_CASE(4)
        default: return 1;
    }
#undef LB_FUSED_CASE
#undef LB_FUSED_ARGS
}

bool lb_sigma_fwd_fused_covers(const lb_plan* plan) {
#ifdef LB_TEST_SWITCHES
    static int disabled = -1;
    if (disabled < 0) {
        const char* e = getenv("LONER_B200_FUSED_FWD");
        disabled = (e && e[0] == '0') ? 1 : 0;
    }
#else
    constexpr int disabled = 0;
#endif
    return !disabled && lb_encode_dense_prefix(plan) >= 1 && lb_encode_single_group(plan) && plan->cfg.n_levels == 16;
}

// returns 1 when the configuration is not covered (the caller then runs the two-kernel path)
int lb_launch_sigma_fwd_fused(const lb_plan* plan, const float* rays, const float* z, const float* x01, const float* params,
                              size_t n_points, int n_samples, float* sigma, float* enc, uint64_t* act, cudaStream_t st) {
    if ((!rays && !x01) || !lb_sigma_fwd_fused_covers(plan)) return 1;
    const lb_level_table T = lb_make_level_table(plan);
    switch (lb_encode_dense_prefix(plan)) {
        case 1: return launch_fused_fwd<1>(plan, T, rays, z, x01, params, n_points, n_samples, sigma, enc, act, st);
        case 2: return launch_fused_fwd<2>(plan, T, rays, z, x01, params, n_points, n_samples, sigma, enc, act, st);
        case 3: return launch_fused_fwd<3>(plan, T, rays, z, x01, params, n_points, n_samples, sigma, enc, act, st);
        case 4: return launch_fused_fwd<4>(plan, T, rays, z, x01, params, n_points, n_samples, sigma, enc, act, st);
        default: return 1;
    }
}

// forward + forward-mode input derivative (sigma_fwd_jvp_fused_kernel); returns 1 when the configuration is not covered
int lb_launch_sigma_fwd_jvp_fused(const lb_plan* plan, const float* rays, const float* z, const float* x01, const float* params,
                                  size_t n_points, int n_samples, float* sigma, float* enc, uint64_t* act, float* dsigma_dx,
                                  cudaStream_t st) {
    if ((!rays && !x01) || !dsigma_dx || !lb_sigma_fwd_fused_covers(plan)) return 1;
    const lb_level_table T = lb_make_level_table(plan);
    switch (lb_encode_dense_prefix(plan)) {
        case 1: return launch_fused_fwd_jvp<1>(plan, T, rays, z, x01, params, n_points, n_samples, sigma, enc, act, dsigma_dx, st);
        case 2: return launch_fused_fwd_jvp<2>(plan, T, rays, z, x01, params, n_points, n_samples, sigma, enc, act, dsigma_dx, st);
        case 3: return launch_fused_fwd_jvp<3>(plan, T, rays, z, x01, params, n_points, n_samples, sigma, enc, act, dsigma_dx, st);
        case 4: return launch_fused_fwd_jvp<4>(plan, T, rays, z, x01, params, n_points, n_samples, sigma, enc, act, dsigma_dx, st);
        default: return 1;
    }
}
