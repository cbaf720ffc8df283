// R2: occupancy-guided + stratified ray sampling, one warp per ray.
//
// Replaces OccGridRaySampler.get_samples / UniformRaySampler.get_samples
// (reference src/models/ray_sampling.py:18-92), OccupancyGridModel.interpolate
// (src/models/model_tcnn.py:126-134) and sample_pdf (src/models/rendering_tcnn.py:19-68).
//
// Bit-exactness contract (oracle/loner_oracle.py): every fp32 operation is issued with explicit
// round-to-nearest intrinsics in the reference's order (no FMA contraction); the pdf normaliser and
// the cdf are accumulated in fp64, which is exact for these operands, so the warp-parallel
// reduction / scan order cannot change a bit; exp(-logit) is the fp64 exp rounded once to fp32.
#include "lb_common.cuh"

namespace {

constexpr int kWarpsPerBlock = 4;

struct RayHead {
    float ox, oy, oz, dx, dy, dz, near, far;
};

__device__ __forceinline__ RayHead load_ray(const float* __restrict__ rays, int r) {
    const float* p = rays + (size_t)r * LB_RAY_STRIDE;
    RayHead h;
    h.ox = __ldg(p + 0); h.oy = __ldg(p + 1); h.oz = __ldg(p + 2);
    h.dx = __ldg(p + 3); h.dy = __ldg(p + 4); h.dz = __ldg(p + 5);
    h.near = __ldg(p + 11); h.far = __ldg(p + 12);
    return h;
}

// near*(1-t) + far*t, each op rounded (ray_sampling.py:61)
__device__ __forceinline__ float z_base(float near, float far, float t) {
    return __fadd_rn(__fmul_rn(near, __fsub_rn(1.0f, t)), __fmul_rn(far, t));
}

// stratified depth k of n (ray_sampling.py:64-73)
__device__ __forceinline__ float z_stratified(const RayHead& h, const float* __restrict__ t, int k, int n,
                                              float perturb, const float* __restrict__ u_row) {
    const float zk = z_base(h.near, h.far, __ldg(t + k));
    if (!(perturb > 0.0f)) return zk;
    float lower = zk, upper = zk;
    if (k > 0) lower = __fmul_rn(0.5f, __fadd_rn(z_base(h.near, h.far, __ldg(t + k - 1)), zk));
    if (k < n - 1) upper = __fmul_rn(0.5f, __fadd_rn(zk, z_base(h.near, h.far, __ldg(t + k + 1))));
    const float pu = __fmul_rn(perturb, __ldg(u_row + k));
    return __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), pu));
}

__device__ __forceinline__ float unnormalize(float c, float size) {
    return __fdiv_rn(__fsub_rn(__fmul_rn(__fadd_rn(c, 1.0f), size), 1.0f), 2.0f);
}

// ATen grid_sampler_3d (bilinear, zeros padding, align_corners=False): corners accumulated in the
// order tnw,tne,tsw,tse,bnw,bne,bsw,bse with separate multiply and add. x -> W (fastest), z -> D.
__device__ __forceinline__ float occ_trilinear(const float* __restrict__ occ, int D, float px, float py, float pz) {
    const float fD = (float)D;
    const float ix = unnormalize(px, fD), iy = unnormalize(py, fD), iz = unnormalize(pz, fD);
    const float x0f = floorf(ix), y0f = floorf(iy), z0f = floorf(iz);
    const float x1f = __fadd_rn(x0f, 1.0f), y1f = __fadd_rn(y0f, 1.0f), z1f = __fadd_rn(z0f, 1.0f);
    const float wx1 = __fsub_rn(ix, x0f), wx0 = __fsub_rn(x1f, ix);
    const float wy1 = __fsub_rn(iy, y0f), wy0 = __fsub_rn(y1f, iy);
    const float wz1 = __fsub_rn(iz, z0f), wz0 = __fsub_rn(z1f, iz);
    const int x0 = (int)x0f, y0 = (int)y0f, z0 = (int)z0f;
    float out = 0.0f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        const int bx = c & 1, by = (c >> 1) & 1, bz = (c >> 2) & 1;
        const int x = x0 + bx, y = y0 + by, z = z0 + bz;
        if (x >= 0 && x < D && y >= 0 && y < D && z >= 0 && z < D) {
            const float w = __fmul_rn(__fmul_rn(bx ? wx1 : wx0, by ? wy1 : wy0), bz ? wz1 : wz0);
            out = __fadd_rn(out, __fmul_rn(__ldg(occ + ((size_t)z * D + y) * D + x), w));
        }
    }
    return out;
}

// ray_sampling.py:79-80 with a correctly rounded exp
__device__ __forceinline__ float occ_prob(float logit) {
    const float e = (float)exp(-(double)logit);
    float p = __fdiv_rn(1.0f, __fadd_rn(1.0f, e));
    p = fminf(fmaxf(p, 0.5f), 1.0f);
    return __fmul_rn(2.0f, __fsub_rn(p, 0.5f));
}

// in-place ascending bitonic sort of n_pow2 floats in shared memory by one warp
__device__ __forceinline__ void warp_bitonic_sort(float* buf, int n_pow2, int lane) {
    for (int k = 2; k <= n_pow2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = lane; i < n_pow2; i += 32) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const float a = buf[i], b = buf[ixj];
                    const bool up = ((i & k) == 0);
                    if ((a > b) == up) { buf[i] = b; buf[ixj] = a; }
                }
            }
            __syncwarp();
        }
    }
}

// The same compare-exchange network on n = 32*R floats held in registers, element i = r*32 + lane: partners at
// distance j < 32 are a lane shuffle, j >= 32 another register of the same lane.  Same swap rule as above.
template <int R>
__device__ __forceinline__ void warp_bitonic_sort_regs(float (&v)[R], int lane) {
#pragma unroll
    for (int k = 2; k <= 32 * R; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j >= 32; j >>= 1) {            // in-lane stages (k >= 64): all indices static
            const int jr = j >> 5, kr = k >> 5;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if ((r & jr) == 0) {
                    const bool up = ((r & kr) == 0);
                    const float a = v[r], b = v[r | jr];
                    if ((a > b) == up) { v[r] = b; v[r | jr] = a; }
                }
            }
        }
#pragma unroll 1
        for (int j = (k >> 1) < 16 ? (k >> 1) : 16; j > 0; j >>= 1) {   // lane-exchange stages
            const bool lower = (lane & j) == 0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const bool up = ((((r << 5) | lane) & k) == 0);
                const float mine = v[r], other = __shfl_xor_sync(0xffffffffu, mine, j);
                const float a = lower ? mine : other, b = lower ? other : mine;
                if ((a > b) == up) v[r] = other;
            }
        }
    }
}

// number of entries of the ascending array a[0, n) that are < x (strict) / <= x
__device__ __forceinline__ int count_less(const float* a, int n, float x) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (a[mid] < x) lo = mid + 1; else hi = mid;
    }
    return lo;
}
__device__ __forceinline__ int count_less_equal(const float* a, int n, float x) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (a[mid] <= x) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// R > 0: the S/2 fine depths stay in registers (S/2 <= 32*R), are sorted there and merged with the coarse depths,
// which the stratified draw already produced in ascending order (checked; any other input takes the full sort).
// R == 0: every depth goes through the shared-memory bitonic sort.
template <int R>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
sample_occ_kernel(const float* __restrict__ rays, const float* __restrict__ occ, int D,
                  const float* __restrict__ t_steps, const float* __restrict__ u1,
                  const float* __restrict__ u2, int n_rays, int S, int SP, float perturb,
                  float* __restrict__ z_out, uint8_t* __restrict__ occ_bits, int64_t* __restrict__ inds_out) {
    extern __shared__ float smem[];
    const int S2 = S >> 1;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ray = blockIdx.x * kWarpsPerBlock + warp;
    if (ray >= n_rays) return;
    // per-warp shared layout: sortbuf[SP] | cdf[S2] | bins[S2]
    float* sortbuf = smem + (size_t)warp * (SP + 2 * S2);
    float* cdf = sortbuf + SP;
    float* bins = cdf + S2;
    float* prob = sortbuf + S2;   // temporary home of the occupancy probabilities (dead before zf lands)

    const RayHead h = load_ray(rays, ray);
    const float* u1r = u1 + (size_t)ray * S2;
    const float* u2r = u2 + (size_t)ray * S2;

    // coarse depths + occupancy probability at the coarse points
    for (int k = lane; k < S2; k += 32) {
        const float zc = z_stratified(h, t_steps, k, S2, perturb, u1r);
        sortbuf[k] = zc;
        const float px = __fadd_rn(h.ox, __fmul_rn(h.dx, zc));
        const float py = __fadd_rn(h.oy, __fmul_rn(h.dy, zc));
        const float pz = __fadd_rn(h.oz, __fmul_rn(h.dz, zc));
        const float p = occ_prob(occ_trilinear(occ, D, px, py, pz));
        prob[k] = p;
        if (occ_bits) occ_bits[(size_t)ray * S2 + k] = (p > 0.0f) ? 1 : 0;
    }
    __syncwarp();

    // pdf normaliser: exact fp64 sum of (p+1e-5) over the S2-2 interior bins, rounded once
    const int n_w = S2 - 2;
    double acc = 0.0;
    for (int j = lane; j < n_w; j += 32) acc += (double)__fadd_rn(prob[j + 1], 1e-5f);
    const float wsum = (float)warp_sum_d(acc);

    // bins = interval mid points (S2-1), cdf = [0, cumsum(pdf)] (S2-1 entries)
    for (int j = lane; j < S2 - 1; j += 32) bins[j] = __fmul_rn(0.5f, __fadd_rn(sortbuf[j], sortbuf[j + 1]));
    if (lane == 0) cdf[0] = 0.0f;
    double carry = 0.0;
    for (int base = 0; base < n_w; base += 32) {
        const int j = base + lane;
        double v = (j < n_w) ? (double)__fdiv_rn(__fadd_rn(prob[j + 1], 1e-5f), wsum) : 0.0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double t = __shfl_up_sync(0xffffffffu, v, o);
            if (lane >= o) v += t;
        }
        v += carry;
        if (j < n_w) cdf[j + 1] = (float)v;
        carry = __shfl_sync(0xffffffffu, v, 31);
    }
    __syncwarp();

    // inverse-CDF draw of the S2 fine depths (searchsorted right=True, rendering_tcnn.py:53-67)
    const int n_cdf = S2 - 1;
    auto draw = [&](int k) -> float {
        const float u = __ldg(u2r + k);
        int lo = 0, hi = n_cdf;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (cdf[mid] <= u) lo = mid + 1; else hi = mid;
        }
        const int below = max(lo - 1, 0), above = min(lo, n_w);
        const float cb = cdf[below], ca = cdf[above];
        float denom = __fsub_rn(ca, cb);
        if (denom < 1e-5f) denom = 1.0f;
        const float bb = bins[below], ba = bins[above];
        if (inds_out) inds_out[(size_t)ray * S2 + k] = (int64_t)lo;
        return __fadd_rn(bb, __fmul_rn(__fdiv_rn(__fsub_rn(u, cb), denom), __fsub_rn(ba, bb)));
    };
    const float kInf = __int_as_float(0x7f800000);
    float* zo = z_out + (size_t)ray * S;
    if constexpr (R > 0) {
        float zf[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int k = r * 32 + lane;
            zf[r] = (k < S2) ? draw(k) : kInf;
        }
        // are the coarse depths ascending?  (always, for the reference's stratified draw with near < far)
        bool ok = true;
        for (int k = lane; k < S2 - 1; k += 32) ok = ok && (sortbuf[k] <= sortbuf[k + 1]);
        ok = __all_sync(0xffffffffu, ok);
        __syncwarp();                                       // all reads of cdf / bins / prob are done
        if (ok) {
            warp_bitonic_sort_regs<R>(zf, lane);
            float* fine = sortbuf + S2;                     // prob's home, dead
            float* merged = cdf;                            // cdf | bins: S contiguous floats, dead
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int k = r * 32 + lane;
                if (k < S2) fine[k] = zf[r];
            }
            __syncwarp();
            // merge by rank: ties place the coarse depth first, so the S positions are a permutation
            for (int k = lane; k < S2; k += 32) {
                const float c = sortbuf[k];
                merged[k + count_less(fine, S2, c)] = c;
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int k = r * 32 + lane;
                if (k < S2) merged[k + count_less_equal(sortbuf, S2, zf[r])] = zf[r];
            }
            __syncwarp();
            for (int i = lane; i < S; i += 32) zo[i] = merged[i];
            return;
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const int k = r * 32 + lane;
            if (k < S2) sortbuf[S2 + k] = zf[r];
        }
    } else {
        for (int k = lane; k < S2; k += 32) {
            const float zf = draw(k);
            sortbuf[S2 + k] = zf;   // overwrites prob[k]: draw() reads cdf / bins only
        }
    }
    for (int i = S + lane; i < SP; i += 32) sortbuf[i] = kInf;
    __syncwarp();

    warp_bitonic_sort(sortbuf, SP, lane);
    for (int i = lane; i < S; i += 32) zo[i] = sortbuf[i];
}

__global__ void sample_uniform_kernel(const float* __restrict__ rays, const float* __restrict__ t_steps,
                                      const float* __restrict__ u, int n_rays, int S, float perturb,
                                      float* __restrict__ z_out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)n_rays * S) return;
    const int ray = (int)(i / S), k = (int)(i % S);
    const RayHead h = load_ray(rays, ray);
    z_out[i] = z_stratified(h, t_steps, k, S, perturb, u + (size_t)ray * S);
}

}  // namespace

extern "C" int lb_sample_occ(const float* rays, const float* occ, int occ_dim, const float* t_steps,
                             const float* u1, const float* u2, int n_rays, int n_samples, float perturb,
                             float* z_out, uint8_t* occ_bits, int64_t* inds, void* stream) {
    LB_REQUIRE(n_rays == 0 || (rays && occ && t_steps && u2 && z_out), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_rays == 0 || perturb <= 0.0f || u1, "perturb > 0 needs u1");
    LB_REQUIRE(n_rays >= 0 && occ_dim > 0, "bad sizes");
    LB_REQUIRE(n_samples >= 8 && (n_samples % 2) == 0 && n_samples <= 8192, "n_samples must be even, in [8, 8192]");
    if (n_rays == 0) return LB_OK;
    int SP = 1;
    while (SP < n_samples) SP <<= 1;
    const size_t smem = (size_t)kWarpsPerBlock * (SP + n_samples) * sizeof(float);
    LB_REQUIRE(smem <= 227 * 1024, "n_samples too large for the shared-memory sort buffer");
    const int blocks = (n_rays + kWarpsPerBlock - 1) / kWarpsPerBlock;
    const int S2 = n_samples / 2;
#define LB_SAMPLE_LAUNCH(R)                                                                                             \
    do {                                                                                                                \
        if (smem > 48 * 1024) {                                                                                         \
            int rc = lb_check_cuda(cudaFuncSetAttribute(sample_occ_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                                        (int)smem), "lb_sample_occ: smem opt-in");                     \
            if (rc) return rc;                                                                                          \
        }                                                                                                               \
        sample_occ_kernel<R><<<blocks, kWarpsPerBlock * 32, smem, (cudaStream_t)stream>>>(                              \
            rays, occ, occ_dim, t_steps, u1, u2, n_rays, n_samples, SP, perturb, z_out, occ_bits, inds);                \
    } while (0)
    if (S2 <= 32) LB_SAMPLE_LAUNCH(1);
    else if (S2 <= 64) LB_SAMPLE_LAUNCH(2);
    else if (S2 <= 128) LB_SAMPLE_LAUNCH(4);
    else if (S2 <= 256) LB_SAMPLE_LAUNCH(8);
    else if (S2 <= 512) LB_SAMPLE_LAUNCH(16);
    else if (S2 <= 1024) LB_SAMPLE_LAUNCH(32);
    else LB_SAMPLE_LAUNCH(0);
#undef LB_SAMPLE_LAUNCH
    LB_LAUNCH_CHECK("lb_sample_occ");
    return LB_OK;
}

extern "C" int lb_sample_uniform(const float* rays, const float* t_steps, const float* u, int n_rays,
                                 int n_samples, float perturb, float* z_out, void* stream) {
    LB_REQUIRE(n_rays == 0 || (rays && t_steps && z_out), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_rays == 0 || perturb <= 0.0f || u, "perturb > 0 needs u");
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1, "bad sizes");
    if (n_rays == 0) return LB_OK;
    const size_t total = (size_t)n_rays * n_samples;
    sample_uniform_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        rays, t_steps, u, n_rays, n_samples, perturb, z_out);
    LB_LAUNCH_CHECK("lb_sample_uniform");
    return LB_OK;
}
