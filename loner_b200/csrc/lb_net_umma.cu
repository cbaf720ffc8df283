// RGB branch, network half: tcnn.Network (FullyFusedMLP, n_in -> 64 x n_hidden -> n_out, ReLU, no bias) forward
// as ONE tcgen05 kernel in which the activations never leave the SM.
//
// Replaces `self._model_intensity = tcnn.Network(...)` / `h_c = self._model_intensity(h_xd)` and the following
// `torch.sigmoid` (reference src/models/nerf_tcnn.py:50-52,86-93; cfg/nerf_config/default_nerf_hash.yaml:8-13).
// Weight layout = tcnn's: W0 [64, n_in], W1.. [64, 64], W_out [16, 64] (output rows padded to 16), row-major, no bias.
//
// Design.  One CTA of 512 threads per SM = two independent groups of 8 warps; each group walks its own 128-point
// tiles (UMMA M = 128); thread (t, half) owns point (TMEM lane) t and one 32-column half of its rows:
//   * the input tile x[128, n_in] is staged with cp.async (coalesced, padded rows); thread t reads row t, splits
//     it exactly into tf32 hi / lo and writes both to tensor memory (tcgen05.st) -- the A operand of every layer
//     lives in TMEM, only the weights (B, hi / lo tiles resident in shared memory for all layers) are read from
//     shared memory by the tensor core;
//   * per layer: K/8 x 3 tcgen05.mma.kind::tf32 (hi*hi + hi*lo + lo*hi, fp32 accumulate, error ~2^-21) into the
//     64-column accumulator, commit -> mbarrier; the epilogue reads the accumulator row (tcgen05.ld), applies ReLU,
//     optionally streams the activation row to global memory for the backward pass, splits it and writes it back
//     to TMEM as the next layer's A operand;
//   * the output layer is one N = 16 MMA group; its epilogue applies the optional sigmoid and stores n_out values.
// While one group runs an epilogue the other one's MMAs occupy the tensor pipe.
#include "lb_common.cuh"
#include "lb_umma.cuh"

namespace {

using namespace lb_umma;

constexpr int kTile = 128;
constexpr int kGroupThreads = 256;      // 8 warps: two column halves per TMEM lane
constexpr int kThreads = 512;
constexpr int kWidth = 64;
constexpr int kOutPad = 16;
constexpr uint32_t kLboW = kWidth * 16 + 16;       // chunk stride of a 64-row weight tile
constexpr uint32_t kLboO = kOutPad * 16 + 16;      // chunk stride of the 16-row output-layer tile
constexpr uint32_t kTmemColsPerGroup = 256;        // accumulator 64 | A hi 64 | A lo 64 | unused 64
constexpr int kMaxHidden = 6;

__device__ __forceinline__ void group_barrier(int group) {
    asm volatile("bar.sync %0, %1;" ::"r"(group + 1), "n"(kGroupThreads) : "memory");
}

struct NetShape {
    int n_in, n_hidden, n_out, out_act;
};

// bytes of one tf32 piece (hi or lo) of all weight tiles
__host__ __device__ inline uint32_t weight_piece_bytes(int n_in, int n_hidden) {
    return (uint32_t)(n_in / 4) * kLboW + (uint32_t)(n_hidden - 1) * (kWidth / 4) * kLboW + (kWidth / 4) * kLboO;
}
// offset of layer l inside a piece; l == n_hidden is the output layer
__device__ __forceinline__ uint32_t weight_tile_off(int n_in, int l) {
    return l == 0 ? 0u : (uint32_t)(n_in / 4) * kLboW + (uint32_t)(l - 1) * (kWidth / 4) * kLboW;
}

// all layers: fp32 [rows, K] row-major -> hi / lo tf32 tiles X[r = out unit][c = k / 4]
__device__ void stage_weights(const float* __restrict__ params, const NetShape s, uint8_t* hi, uint8_t* lo) {
    size_t src = 0;
    for (int l = 0; l <= s.n_hidden; ++l) {
        const int rows = l == s.n_hidden ? kOutPad : kWidth;
        const int k = l == 0 ? s.n_in : kWidth;
        const uint32_t lbo = l == s.n_hidden ? kLboO : kLboW;
        const uint32_t off = weight_tile_off(s.n_in, l);
        const int chunks = k / 4;
        for (int it = threadIdx.x; it < rows * chunks; it += kThreads) {
            const int j = it / chunks, c = it - j * chunks;
            const float4 v = __ldg(reinterpret_cast<const float4*>(params + src + (size_t)j * k + c * 4));
            uint4 a, b;
            split_tf32(v, a, b);
            *reinterpret_cast<uint4*>(hi + off + tile_off(j, c, lbo)) = a;
            *reinterpret_cast<uint4*>(lo + off + tile_off(j, c, lbo)) = b;
        }
        src += (size_t)rows * k;
    }
}

template <int kIn>
__device__ __forceinline__ void stage_input(const float* __restrict__ x, size_t tile, size_t n_points, uint32_t dst,
                                            int gtid) {
    constexpr int kChunks = kIn / 4;
    constexpr uint32_t kRow = kIn * 4 + 16;
    const float4* base = reinterpret_cast<const float4*>(x) + tile * (size_t)(kTile * kChunks);
    const size_t first = tile * (size_t)kTile;
#pragma unroll
    for (int i = 0; i < (kTile * kChunks) / kGroupThreads; ++i) {
        const int idx = i * kGroupThreads + gtid;
        const int row = idx / kChunks, q = idx - row * kChunks;
        const bool ok = first + row < n_points;
        cp_async16(dst + row * kRow + q * 16, ok ? (const void*)(base + idx) : (const void*)x, ok ? 16u : 0u);
    }
}

// 16 fp32 values -> tf32 hi / lo -> 16 TMEM columns of this thread's lane
__device__ __forceinline__ void put_a16(const float* v, uint32_t t_hi, uint32_t t_lo) {
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        hi[i] = __float_as_uint(v[i]) & 0xffffe000u;
        lo[i] = __float_as_uint(v[i] - __uint_as_float(hi[i]));
    }
    tmem_st16(t_hi, hi);
    tmem_st16(t_lo, lo);
}

template <int kIn>
__global__ void __launch_bounds__(kThreads, 1)
net_fwd_umma_kernel(const float* __restrict__ params, const float* __restrict__ x, size_t n_points, NetShape shape,
                    float* __restrict__ out, float* __restrict__ acts) {
    extern __shared__ __align__(128) uint8_t sm[];
    constexpr uint32_t kRow = kIn * 4 + 16;
    constexpr uint32_t kStage = kTile * kRow;
    // group = one 128-point tile stream; inside a group, thread (lane_t, half) owns TMEM lane lane_t and the column
    // half `half` of every 64-wide row (warps w and w + 4 of a group reach the same 32 TMEM lanes)
    const int tid = threadIdx.x, group = tid / kGroupThreads, gtid = tid % kGroupThreads, warp = tid >> 5;
    const int lane_t = gtid & 127, half = gtid >> 7;
    const uint32_t piece = weight_piece_bytes(kIn, shape.n_hidden);
    // carve-up: [hi piece][lo piece][stage g0][stage g1][2 mbarriers][tmem slot]
    uint8_t* w_hi = sm;
    uint8_t* w_lo = sm + piece;
    uint8_t* stage = sm + 2 * piece + group * kStage;
    const uint32_t bar = smem_u32(sm + 2 * piece + 2 * kStage + group * 8);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + 2 * piece + 2 * kStage + 16);

    const size_t n_tiles = (n_points + kTile - 1) / kTile;
    const size_t stride = (size_t)gridDim.x * 2;
    size_t tile = (size_t)blockIdx.x * 2 + group;
    if (tile < n_tiles) stage_input<kIn>(x, tile, n_points, smem_u32(stage), gtid);
    cp_async_commit();

    if (warp == 0) tmem_alloc(smem_u32(tmem_slot), 512);
    if (gtid == 32) mbar_init(bar, 1);
    stage_weights(params, shape, w_hi, w_lo);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot + group * kTmemColsPerGroup;
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t t_acc = tmem, t_ahi = tmem + 64, t_alo = tmem + 128;
    const uint32_t whi = smem_u32(w_hi), wlo = smem_u32(w_lo);
    constexpr uint32_t idesc_h = umma_idesc_tf32(128, kWidth);
    constexpr uint32_t idesc_o = umma_idesc_tf32(128, kOutPad);

    uint32_t parity = 0;
    for (; tile < n_tiles; tile += stride) {
        cp_async_wait<0>();
        group_barrier(group);                                          // the staged input tile is complete
        {
            const uint8_t* row = stage + lane_t * kRow;
#pragma unroll
            for (int c = 0; c < kIn / 16; ++c) {
                if ((c & 1) != half) continue;                          // 16-column chunks alternate between the halves
                float v[16];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 f = *reinterpret_cast<const float4*>(row + (c * 4 + q) * 16);
                    v[4 * q] = f.x; v[4 * q + 1] = f.y; v[4 * q + 2] = f.z; v[4 * q + 3] = f.w;
                }
                put_a16(v, t_ahi + lane_off + c * 16, t_alo + lane_off + c * 16);
            }
            tmem_st_wait();
        }
        tc_fence_before();
        group_barrier(group);                                          // A is in TMEM; the staging tile is free
        if (tile + stride < n_tiles) stage_input<kIn>(x, tile + stride, n_points, smem_u32(stage), gtid);
        cp_async_commit();

        const size_t p = tile * (size_t)kTile + lane_t;
        for (int l = 0; l <= shape.n_hidden; ++l) {
            const bool last = l == shape.n_hidden;
            if (gtid == 0) {
                tc_fence_after();
                const int ksteps = (l == 0 ? kIn : kWidth) / 8;
                const uint32_t lbo = last ? kLboO : kLboW;
                const uint32_t off = weight_tile_off(kIn, l);
                const uint32_t idesc = last ? idesc_o : idesc_h;
                for (int ks = 0; ks < ksteps; ++ks) {
                    const uint64_t dbh = umma_desc(whi + off + ks * 2 * lbo, lbo, 128);
                    const uint64_t dbl = umma_desc(wlo + off + ks * 2 * lbo, lbo, 128);
                    umma_tf32_ts(t_acc, t_ahi + ks * 8, dbh, idesc, ks > 0);
                    umma_tf32_ts(t_acc, t_ahi + ks * 8, dbl, idesc, 1);
                    umma_tf32_ts(t_acc, t_alo + ks * 8, dbh, idesc, 1);
                }
                umma_commit(bar);
            }
            mbar_wait(bar, parity);
            parity ^= 1;
            tc_fence_after();
            if (!last) {
                // hidden layer epilogue on this thread's 32 columns: ReLU, optional activation stream-out (tile-transposed
                // layout [layer][tile][16-byte chunk][row]: a warp store covers 512 contiguous bytes), next A operand
                uint32_t r[32];
                tmem_ld32(t_acc + lane_off + half * 32, r);
                tmem_ld_wait();
                float4* act_tile = acts ? reinterpret_cast<float4*>(acts) + ((size_t)l * n_tiles + tile) * (kTile * 16) : nullptr;
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    float v[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) v[i] = fmaxf(__uint_as_float(r[c * 16 + i]), 0.0f);
                    if (act_tile) {
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            act_tile[(half * 8 + c * 4 + q) * kTile + lane_t] =
                                make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
                    }
                    put_a16(v, t_ahi + lane_off + half * 32 + c * 16, t_alo + lane_off + half * 32 + c * 16);
                }
                tmem_st_wait();
                tc_fence_before();
                group_barrier(group);
            } else {
                if (half == 0) {
                    uint32_t r[16];
                    tmem_ld16(t_acc + lane_off, r);
                    tmem_ld_wait();
                    if (p < n_points) {
#pragma unroll
                        for (int i = 0; i < kOutPad; ++i) {
                            if (i < shape.n_out) {
                                float o = __uint_as_float(r[i]);
                                if (shape.out_act == 1) o = 1.0f / (1.0f + expf(-o));
                                out[p * shape.n_out + i] = o;
                            }
                        }
                    }
                }
                tc_fence_before();      // orders the accumulator reads before the next tile's first MMA (after the barriers)
            }
        }
    }
    cp_async_wait<0>();
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(*tmem_slot, 512);
}

int device_sm_count() {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
            sms = 148;
    }
    return sms;
}

template <int kIn>
int launch_fwd(const float* params, const float* x, size_t n_points, NetShape shape, float* out, float* acts,
               cudaStream_t st) {
    const uint32_t smem = 2 * weight_piece_bytes(kIn, shape.n_hidden) + 2 * kTile * (kIn * 4 + 16) + 32;
    if (smem > 227 * 1024) {
        lb_set_error("lb_net_fwd: %d hidden layers of width 64 do not fit in shared memory", shape.n_hidden);
        return LB_ERR_UNSUPPORTED;
    }
    int rc = lb_check_cuda(cudaFuncSetAttribute(net_fwd_umma_kernel<kIn>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                           "net_fwd_umma smem opt-in");
    if (rc) return rc;
    const size_t tiles = (n_points + kTile - 1) / kTile;
    size_t grid = (tiles + 1) / 2;
    if (grid > (size_t)device_sm_count()) grid = device_sm_count();
    net_fwd_umma_kernel<kIn><<<(int)grid, kThreads, smem, st>>>(params, x, n_points, shape, out, acts);
    return lb_check_cuda(cudaGetLastError(), "net_fwd_umma launch");
}

// ------------------------------------------------------------------------------------------
// backward on the tensor cores.  Per 128-point tile, from the output layer down:
//   U_l[j,k] += sum_p dZ_l[p,j] Hprev[p,k]      A = dZ_l^T (MN-major view),  B = Hprev (MN-major view)   M = 64
//   dHprev[p,k] = sum_j dZ_l[p,j] W_l[j,k]      A = dZ_l   (K-major view),   B = W_l   (MN-major view)   M = 128
//   dZ_{l-1} = dHprev * [Hprev > 0]
// Operands are exact two-piece bf16 splits (x = b1 + b2 up to 2^-18), three products per K step, fp32 accumulate.
// Thread (t, half) owns row t / 32 columns: its dZ half-row lives in registers between layers, its Hprev half-row
// is one coalesced read of the tile-transposed activation buffer; both are written as bf16 pieces into
// shared-memory operand tiles (canonical no-swizzle layout, row-per-thread stores are conflict free).  The weight
// pieces of all layers stay resident in shared memory; the weight-gradient accumulators U_l stay in tensor memory
// (64 columns per layer, M = 64 layout) for the CTA's lifetime and are flushed with one atomicAdd per entry.
// The output layer is treated as a 64-wide layer whose rows >= n_out are zero.
// ------------------------------------------------------------------------------------------
constexpr uint32_t kLboP = kTile * 16 + 16;          // chunk stride of a 128-row bf16 operand tile
constexpr uint32_t kPiece = 8 * kLboP;               // [128 x 64] bf16 piece
constexpr int kBwdThreads = 256;

__host__ __device__ inline uint32_t bwd_weight_piece_bytes(int n_in, int n_hidden) {
    return (uint32_t)(n_in / 8) * kLboW + (uint32_t)n_hidden * (kWidth / 8) * kLboW;     // layers 0 .. n_hidden
}
__device__ __forceinline__ uint32_t bwd_weight_off(int n_in, int l) {
    return l == 0 ? 0u : (uint32_t)(n_in / 8) * kLboW + (uint32_t)(l - 1) * (kWidth / 8) * kLboW;
}

// 8 fp32 -> bf16 pieces -> one 16-byte chunk of each piece tile
__device__ __forceinline__ void put_chunk(const float* v, uint8_t* p1, uint8_t* p2, uint32_t off) {
    uint32_t b1[4], b2[4];
    split8<false>(v, b1, b2, nullptr);
    *reinterpret_cast<uint4*>(p1 + off) = make_uint4(b1[0], b1[1], b1[2], b1[3]);
    *reinterpret_cast<uint4*>(p2 + off) = make_uint4(b2[0], b2[1], b2[2], b2[3]);
}

// this thread's 32-column half of row t of an activation tile (tile-transposed layout: coalesced across t)
__device__ __forceinline__ void load_act_half(const float* __restrict__ acts, size_t layer, size_t n_tiles, size_t tile,
                                              int half, int t, float* h) {
    const float4* src = reinterpret_cast<const float4*>(acts) + (layer * n_tiles + tile) * (size_t)(kTile * 16);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
        const float4 v = __ldcs(src + (half * 8 + q) * kTile + t);
        h[4 * q] = v.x; h[4 * q + 1] = v.y; h[4 * q + 2] = v.z; h[4 * q + 3] = v.w;
    }
}

template <int kIn>
__global__ void __launch_bounds__(kBwdThreads, 1)
net_bwd_umma_kernel(const float* __restrict__ params, const float* __restrict__ x, const float* __restrict__ acts,
                    const float* __restrict__ out, const float* __restrict__ d_out, size_t n_points, NetShape shape,
                    float* __restrict__ grad_params, float* __restrict__ d_x) {
    extern __shared__ __align__(128) uint8_t sm[];
    constexpr uint32_t kRow = kIn * 4 + 16;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int t = tid & 127, half = tid >> 7;
    const int nh = shape.n_hidden;
    const uint32_t wpiece = bwd_weight_piece_bytes(kIn, nh);
    uint8_t* dz1 = sm;
    uint8_t* dz2 = dz1 + kPiece;
    uint8_t* h1 = dz2 + kPiece;
    uint8_t* h2 = h1 + kPiece;
    uint8_t* w1 = h2 + kPiece;
    uint8_t* w2 = w1 + wpiece;
    uint8_t* xst = w2 + wpiece;
    const uint32_t bar = smem_u32(xst + kTile * kRow);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(xst + kTile * kRow + 8);

    const size_t n_tiles = (n_points + kTile - 1) / kTile;
    size_t tile = blockIdx.x;
    {   // first x tile (256 threads)
        constexpr int kChunks = kIn / 4;
        const float4* base = reinterpret_cast<const float4*>(x) + tile * (size_t)(kTile * kChunks);
        for (int idx = tid; idx < kTile * kChunks; idx += kBwdThreads) {
            const int row = idx / kChunks, q = idx - row * kChunks;
            const bool ok = tile < n_tiles && tile * (size_t)kTile + row < n_points;
            cp_async16(smem_u32(xst) + row * kRow + q * 16, ok ? (const void*)(base + idx) : (const void*)x, ok ? 16u : 0u);
        }
        cp_async_commit();
    }
    if (warp == 0) tmem_alloc(smem_u32(tmem_slot), 512);
    if (tid == 32) mbar_init(bar, 1);
    // weight pieces of all layers: X[r = out unit j][c = k / 8]; output layer rows >= 16 are zero
    {
        size_t src = 0;
        for (int l = 0; l <= nh; ++l) {
            const int K = l == 0 ? kIn : kWidth;
            const int rows_real = l == nh ? kOutPad : kWidth;
            const int chunks = K / 8;
            const uint32_t off = bwd_weight_off(kIn, l);
            for (int it = tid; it < kWidth * chunks; it += kBwdThreads) {
                const int j = it / chunks, c = it - j * chunks;
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = j < rows_real ? __ldg(params + src + (size_t)j * K + c * 8 + e) : 0.0f;
                put_chunk(v, w1, w2, off + tile_off(j, c, kLboW));
            }
            src += (size_t)rows_real * K;
        }
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t t_dh = tmem;                                    // dHprev [128 x 64]
    const uint32_t lane_off = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t a1 = smem_u32(dz1), a2 = smem_u32(dz2), b1s = smem_u32(h1), b2s = smem_u32(h2);
    const uint32_t ws1 = smem_u32(w1), ws2 = smem_u32(w2);
    const bool need_dx = d_x != nullptr;

    uint32_t parity = 0;
    bool first = true;
    for (; tile < n_tiles; tile += gridDim.x) {
        const size_t p = tile * (size_t)kTile + t;
        const bool valid = p < n_points;
        float dz[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) dz[i] = 0.0f;
        if (half == 0 && valid) {                                     // output layer: dZ = d_out (* o (1 - o))
#pragma unroll
            for (int c = 0; c < kOutPad; ++c) {
                if (c < shape.n_out) {
                    float g = __ldg(d_out + p * shape.n_out + c);
                    if (shape.out_act == 1) {
                        const float o = __ldg(out + p * shape.n_out + c);
                        g *= o * (1.0f - o);
                    }
                    dz[c] = g;
                }
            }
        }
        float hp_next[32];
        load_act_half(acts, (size_t)(nh - 1), n_tiles, tile, half, t, hp_next);
        for (int l = nh; l >= 0; --l) {
            const int K = l == 0 ? kIn : kWidth;
            // dZ_l half-row -> operand tiles
#pragma unroll
            for (int c = 0; c < 4; ++c) put_chunk(dz + 8 * c, dz1, dz2, tile_off(t, half * 4 + c, kLboP));
            // Hprev half-row: activations of layer l-1 (prefetched from the tile-transposed buffer while the previous
            // layer's MMAs ran) or the staged input row
            float hp[32];
            if (l > 0) {
#pragma unroll
                for (int i = 0; i < 32; ++i) hp[i] = hp_next[i];
            } else {
                cp_async_wait<0>();
                __syncthreads();                                           // every thread's copies of the x tile landed
                const float* row = reinterpret_cast<const float*>(xst + t * kRow);
#pragma unroll
                for (int i = 0; i < 32; ++i) hp[i] = (half * 32 + i < kIn) ? row[half * 32 + i] : 0.0f;
            }
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (half * 32 + 8 * c < K) put_chunk(hp + 8 * c, h1, h2, tile_off(t, half * 4 + c, kLboP));
            fence_async_smem();
            tc_fence_before();
            __syncthreads();
            if (tid == 0) {
                tc_fence_after();
                const uint32_t t_u = tmem + 64 + 64 * l;
                const uint32_t idesc_u = umma_idesc(64, K, 1, 1);
                if (grad_params) {
#pragma unroll
                    for (int ks = 0; ks < 8; ++ks) {                   // contraction over the 128 points, 16 per step
                        const uint64_t da1 = umma_desc(a1 + ks * 256, 128, kLboP), da2 = umma_desc(a2 + ks * 256, 128, kLboP);
                        const uint64_t db1 = umma_desc(b1s + ks * 256, 128, kLboP), db2 = umma_desc(b2s + ks * 256, 128, kLboP);
                        umma_bf16(t_u, da1, db1, idesc_u, (!first || ks > 0) ? 1u : 0u);
                        umma_bf16(t_u, da1, db2, idesc_u, 1);
                        umma_bf16(t_u, da2, db1, idesc_u, 1);
                    }
                }
                if (l > 0 || need_dx) {
                    const uint32_t idesc_h = umma_idesc(128, K, 0, 1);
                    const uint32_t woff = bwd_weight_off(kIn, l);
                    const int ksteps = l == nh ? 1 : 4;              // contraction over the layer's outputs j (16 per step)
                    for (int ks = 0; ks < ksteps; ++ks) {
                        const uint64_t da1 = umma_desc(a1 + ks * 2 * kLboP, kLboP, 128), da2 = umma_desc(a2 + ks * 2 * kLboP, kLboP, 128);
                        const uint64_t db1 = umma_desc(ws1 + woff + ks * 256, 128, kLboW), db2 = umma_desc(ws2 + woff + ks * 256, 128, kLboW);
                        umma_bf16(t_dh, da1, db1, idesc_h, ks > 0);
                        umma_bf16(t_dh, da1, db2, idesc_h, 1);
                        umma_bf16(t_dh, da2, db1, idesc_h, 1);
                    }
                }
                umma_commit(bar);
            }
            if (l > 1) load_act_half(acts, (size_t)(l - 2), n_tiles, tile, half, t, hp_next);   // in flight during the MMAs
            mbar_wait(bar, parity);
            parity ^= 1;
            tc_fence_after();
            if (l > 0) {
                uint32_t r[32];
                tmem_ld32(t_dh + lane_off + half * 32, r);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) dz[i] = hp[i] > 0.0f ? __uint_as_float(r[i]) : 0.0f;
            } else {
                if (need_dx && half * 32 < kIn) {
                    uint32_t r[32];
                    tmem_ld32(t_dh + lane_off + half * 32, r);
                    tmem_ld_wait();
                    if (valid) {
#pragma unroll
                        for (int q = 0; q < 8; ++q)
                            if (half * 32 + 4 * q < kIn)
                                *reinterpret_cast<float4*>(d_x + p * kIn + half * 32 + 4 * q) =
                                    make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                                                __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
                    }
                }
                // the staged x rows have been consumed (before this layer's barrier): prefetch the next tile's
                const size_t next = tile + gridDim.x;
                constexpr int kChunks = kIn / 4;
                if (next < n_tiles) {
                    const float4* base = reinterpret_cast<const float4*>(x) + next * (size_t)(kTile * kChunks);
                    for (int idx = tid; idx < kTile * kChunks; idx += kBwdThreads) {
                        const int row = idx / kChunks, q = idx - row * kChunks;
                        const bool ok = next * (size_t)kTile + row < n_points;
                        cp_async16(smem_u32(xst) + row * kRow + q * 16, ok ? (const void*)(base + idx) : (const void*)x, ok ? 16u : 0u);
                    }
                }
                cp_async_commit();
            }
            tc_fence_before();
        }
        first = false;
    }
    cp_async_wait<0>();
    // flush the weight gradients: U_l row j sits in TMEM lane (j % 16) + 32 * (j / 16)
    if (grad_params && !first) {
        tc_fence_after();
        size_t off = 0;
        for (int l = 0; l <= nh; ++l) {
            const int K = l == 0 ? kIn : kWidth;
            const int rows_real = l == nh ? shape.n_out : kWidth;
            uint32_t r[32];
            tmem_ld32(tmem + 64 + 64 * l + lane_off + half * 32, r);
            tmem_ld_wait();
            const int j = (warp & 3) * 16 + lane;
            if (lane < 16 && j < rows_real) {
#pragma unroll
                for (int k = 0; k < 32; ++k)
                    if (half * 32 + k < K) atomicAdd(grad_params + off + (size_t)j * K + half * 32 + k, __uint_as_float(r[k]));
            }
            off += (size_t)(l == nh ? kOutPad : kWidth) * K;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int kIn>
int launch_bwd(const float* params, const float* x, const float* acts, const float* out, const float* d_out,
               size_t n_points, NetShape shape, float* grad_params, float* d_x, cudaStream_t st) {
    const uint32_t smem = 4 * kPiece + 2 * bwd_weight_piece_bytes(kIn, shape.n_hidden) + kTile * (kIn * 4 + 16) + 32;
    if (smem > 227 * 1024) return 1;          // caller falls back to the CUDA-core kernel
    int rc = lb_check_cuda(cudaFuncSetAttribute(net_bwd_umma_kernel<kIn>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                           "net_bwd_umma smem opt-in");
    if (rc) return rc;
    const size_t tiles = (n_points + kTile - 1) / kTile;
    const int grid = (int)(tiles < (size_t)device_sm_count() ? tiles : (size_t)device_sm_count());
    net_bwd_umma_kernel<kIn><<<grid, kBwdThreads, smem, st>>>(params, x, acts, out, d_out, n_points, shape, grad_params, d_x);
    return lb_check_cuda(cudaGetLastError(), "net_bwd_umma launch");
}

}  // namespace

// returns 0 on success, < 0 on error, 1 when the configuration does not fit (caller uses the CUDA-core kernel)
int lb_launch_net_bwd_umma(const float* x, size_t n_points, int n_in, int n_hidden, int n_out, int out_act, const float* params,
                           const float* acts, const float* out, const float* d_out, float* grad_params, float* d_x,
                           cudaStream_t st) {
    const NetShape shape{n_in, n_hidden, n_out, out_act};
    switch (n_in) {
        case 16: return launch_bwd<16>(params, x, acts, out, d_out, n_points, shape, grad_params, d_x, st);
        case 32: return launch_bwd<32>(params, x, acts, out, d_out, n_points, shape, grad_params, d_x, st);
        case 48: return launch_bwd<48>(params, x, acts, out, d_out, n_points, shape, grad_params, d_x, st);
        default: return launch_bwd<64>(params, x, acts, out, d_out, n_points, shape, grad_params, d_x, st);
    }
}

// tcnn.Network forward (reference src/models/nerf_tcnn.py:50-52,89-91).  x [P, n_in] (n_in already padded to a multiple
// of 16), params = tcnn layout, out [P, n_out]; acts [n_hidden, P, 64] receives the post-ReLU activations when the
// backward pass will follow (NULL for inference); out_act: 0 = none, 1 = sigmoid (nerf_tcnn.py:93 fused).
extern "C" int lb_net_fwd(const float* x, int64_t n_points, int n_in, int n_hidden, int n_out, int out_act,
                          const float* params, float* out, float* acts, void* stream) {
    LB_REQUIRE(n_points == 0 || (x && params && out), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_points >= 0, "negative size");
    LB_REQUIRE(n_in >= 16 && n_in <= 64 && n_in % 16 == 0, "n_in must be 16, 32, 48 or 64 (pad the input to a multiple of 16)");
    LB_REQUIRE(n_hidden >= 1 && n_hidden <= kMaxHidden, "n_hidden_layers must be in [1, 6]");
    LB_REQUIRE(n_out >= 1 && n_out <= kOutPad, "n_out must be in [1, 16]");
    LB_REQUIRE(out_act == 0 || out_act == 1, "out_act: 0 = none, 1 = sigmoid");
    if (n_points == 0) return LB_OK;
    const NetShape shape{n_in, n_hidden, n_out, out_act};
    cudaStream_t st = (cudaStream_t)stream;
    switch (n_in) {
        case 16: return launch_fwd<16>(params, x, (size_t)n_points, shape, out, acts, st);
        case 32: return launch_fwd<32>(params, x, (size_t)n_points, shape, out, acts, st);
        case 48: return launch_fwd<48>(params, x, (size_t)n_points, shape, out, acts, st);
        default: return launch_fwd<64>(params, x, (size_t)n_points, shape, out, acts, st);
    }
}
