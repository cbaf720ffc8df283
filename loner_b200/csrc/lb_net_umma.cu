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

// 32 lanes x 16 consecutive columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
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

}  // namespace

// tcnn.Network forward (reference src/models/nerf_tcnn.py:50-52,89-91).  x [P, n_in] (n_in already padded to a multiple
// of 16), params = tcnn layout, out [P, n_out]; acts [n_hidden, P, 64] receives the post-ReLU activations when the
// backward pass will follow (NULL for inference); out_act: 0 = none, 1 = sigmoid (nerf_tcnn.py:93 fused).
extern "C" int lb_net_fwd(const float* x, int64_t n_points, int n_in, int n_hidden, int n_out, int out_act,
                          const float* params, float* out, float* acts, void* stream) {
    LB_REQUIRE(x && params && out, "null pointer");
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
