// R4: multi-resolution hash-grid encoding + 32->64->1 sigma MLP, forward and backward.
//
// Replaces tcnn.NetworkWithInputEncoding / tcnn.Encoding as used by DecoupledNeRF
// (reference src/models/nerf_tcnn.py:35-40,63-72), fused with the sample-point construction
// xyz = o + d*z (src/models/rendering_tcnn.py:390) and pos01 = (xyz+1)/2 (nerf_tcnn.py:63).
// Grid semantics follow the published tiny-cuda-nn algorithm (see oracle/tinycudann/__init__.py).
//
// Thread mapping: one thread per sample point, points ordered [ray, sample] so a warp holds 32
// consecutive samples of one ray: on coarse levels its lanes gather the same few cells (served as a
// handful of 32 B sectors), on fine levels each float2 corner is its own sector.
#include "lb_common.cuh"

namespace {

constexpr int kThreads = 128;

struct Point {
    float x, y, z;   // position in [0,1]^3
};

template <bool kFromRays>
__device__ __forceinline__ Point load_point(const float* __restrict__ rays, const float* __restrict__ zv,
                                            const float* __restrict__ x01, size_t p, int n_samples, float* z_out) {
    Point q;
    if (kFromRays) {
        const size_t r = p / (size_t)n_samples;
        const float* ray = rays + r * LB_RAY_STRIDE;
        const float z = __ldg(zv + p);
        // same roundings as the reference's torch ops (o + d*z, then (xyz+1)/2): at the finest level one
        // ulp of x01 is 3 % of a cell, so the sample position must match bit for bit
        q.x = __fmul_rn(__fadd_rn(__fadd_rn(__ldg(ray + 0), __fmul_rn(__ldg(ray + 3), z)), 1.0f), 0.5f);
        q.y = __fmul_rn(__fadd_rn(__fadd_rn(__ldg(ray + 1), __fmul_rn(__ldg(ray + 4), z)), 1.0f), 0.5f);
        q.z = __fmul_rn(__fadd_rn(__fadd_rn(__ldg(ray + 2), __fmul_rn(__ldg(ray + 5), z)), 1.0f), 0.5f);
        if (z_out) *z_out = z;
    } else {
        q.x = __ldg(x01 + 3 * p + 0);
        q.y = __ldg(x01 + 3 * p + 1);
        q.z = __ldg(x01 + 3 * p + 2);
    }
    return q;
}

struct Cell {
    uint32_t cx, cy, cz;
    float fx, fy, fz;
};

__device__ __forceinline__ Cell locate(const lb_level& L, const Point& q) {
    Cell c;
    const float px = fmaf(L.scale, q.x, 0.5f), py = fmaf(L.scale, q.y, 0.5f), pz = fmaf(L.scale, q.z, 0.5f);
    const float flx = floorf(px), fly = floorf(py), flz = floorf(pz);
    c.cx = (uint32_t)(int)flx; c.cy = (uint32_t)(int)fly; c.cz = (uint32_t)(int)flz;
    c.fx = px - flx; c.fy = py - fly; c.fz = pz - flz;
    return c;
}

__device__ __forceinline__ float corner_weight(const Cell& c, int k) {
    const float wx = (k & 1) ? c.fx : 1.0f - c.fx;
    const float wy = (k & 2) ? c.fy : 1.0f - c.fy;
    const float wz = (k & 4) ? c.fz : 1.0f - c.fz;
    return wx * wy * wz;
}

__device__ __forceinline__ uint32_t corner_index(const lb_level& L, const Cell& c, int k) {
    return L.offset + lb_grid_index(L, c.cx + (k & 1), c.cy + ((k >> 1) & 1), c.cz + ((k >> 2) & 1));
}

// e[2l], e[2l+1] = sum_c w_c * table[off_l + idx_c]
__device__ __forceinline__ void encode_point(const lb_level_table& T, const float2* __restrict__ table,
                                             const Point& q, float* e) {
#pragma unroll
    for (int l = 0; l < 16; ++l) {
        if (l < T.n_levels) {
            const lb_level& L = T.lv[l];
            const Cell c = locate(L, q);
            float2 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = __ldg(table + corner_index(L, c, k));
            float a = 0.0f, b = 0.0f;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float w = corner_weight(c, k);
                a = fmaf(w, v[k].x, a);
                b = fmaf(w, v[k].y, b);
            }
            e[2 * l] = a;
            e[2 * l + 1] = b;
        } else {
            e[2 * l] = 0.0f;
            e[2 * l + 1] = 0.0f;
        }
    }
}

// hidden layer: h_j = relu(sum_k W1[j][k] e_k) with W1 in shared memory (row-major [64][32])
__device__ __forceinline__ float hidden_unit(const float* __restrict__ w1s, int j, const float* e) {
    const float4* row = reinterpret_cast<const float4*>(w1s + j * LB_MLP_IN);
    float acc = 0.0f;
#pragma unroll
    for (int k4 = 0; k4 < LB_MLP_IN / 4; ++k4) {
        const float4 w = row[k4];
        acc = fmaf(w.x, e[4 * k4 + 0], acc);
        acc = fmaf(w.y, e[4 * k4 + 1], acc);
        acc = fmaf(w.z, e[4 * k4 + 2], acc);
        acc = fmaf(w.w, e[4 * k4 + 3], acc);
    }
    return acc;
}

__device__ __forceinline__ void load_mlp_to_smem(const float* __restrict__ mlp, float* w1s, float* w2s) {
    for (int i = threadIdx.x; i < LB_MLP_W * LB_MLP_IN; i += blockDim.x) w1s[i] = __ldg(mlp + i);
    for (int i = threadIdx.x; i < LB_MLP_W; i += blockDim.x) w2s[i] = __ldg(mlp + LB_MLP_W * LB_MLP_IN + i);
}

// ------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------
template <bool kFromRays>
__global__ void __launch_bounds__(kThreads)
sigma_fwd_kernel(lb_level_table T, const float* __restrict__ rays, const float* __restrict__ zv,
                 const float* __restrict__ x01, const float* __restrict__ params, size_t n_points,
                 int n_samples, float* __restrict__ sigma, float* __restrict__ enc_save) {
    __shared__ __align__(16) float w1s[LB_MLP_W * LB_MLP_IN];
    __shared__ float w2s[LB_MLP_W];
    load_mlp_to_smem(params, w1s, w2s);
    __syncthreads();
    const float2* table = reinterpret_cast<const float2*>(params + LB_MLP_PARAMS);
    for (size_t p = (size_t)blockIdx.x * kThreads + threadIdx.x; p < n_points; p += (size_t)gridDim.x * kThreads) {
        const Point q = load_point<kFromRays>(rays, zv, x01, p, n_samples, nullptr);
        float e[LB_MLP_IN];
        encode_point(T, table, q, e);
        if (enc_save) {
            float4* dst = reinterpret_cast<float4*>(enc_save + p * LB_MLP_IN);
#pragma unroll
            for (int k4 = 0; k4 < LB_MLP_IN / 4; ++k4) dst[k4] = make_float4(e[4 * k4], e[4 * k4 + 1], e[4 * k4 + 2], e[4 * k4 + 3]);
        }
        float s = 0.0f;
#pragma unroll 4
        for (int j = 0; j < LB_MLP_W; ++j) s = fmaf(w2s[j], fmaxf(hidden_unit(w1s, j, e), 0.0f), s);
        sigma[p] = s;
    }
}

__global__ void __launch_bounds__(kThreads)
encode_fwd_kernel(lb_level_table T, const float* __restrict__ x01, const float2* __restrict__ table,
                  size_t n_points, float* __restrict__ enc) {
    const int width = T.n_levels * LB_FEATS;
    for (size_t p = (size_t)blockIdx.x * kThreads + threadIdx.x; p < n_points; p += (size_t)gridDim.x * kThreads) {
        const Point q = load_point<false>(nullptr, nullptr, x01, p, 1, nullptr);
        float e[LB_MLP_IN];
        encode_point(T, table, q, e);
        float* dst = enc + p * width;
#pragma unroll
        for (int k = 0; k < LB_MLP_IN; ++k)
            if (k < width) dst[k] = e[k];
    }
}

// ------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------
constexpr int kLd = kThreads + 1;   // padded leading dimension of the [row][point] staging tiles

// scatter de into the table gradient and (optionally) accumulate dL/dx01
template <bool kInputGrad>
__device__ __forceinline__ void scatter_point(const lb_level_table& T, const float2* __restrict__ table,
                                              float* __restrict__ grad_table, const Point& q, const float* de,
                                              float* gx, float* gy, float* gz) {
#pragma unroll
    for (int l = 0; l < 16; ++l) {
        if (l < T.n_levels) {
            const lb_level& L = T.lv[l];
            const Cell c = locate(L, q);
            const float d0 = de[2 * l], d1 = de[2 * l + 1];
            uint32_t idx[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) idx[k] = corner_index(L, c, k);
            if (kInputGrad) {
                float val[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float2 v = __ldg(table + idx[k]);
                    val[k] = fmaf(v.x, d0, v.y * d1);
                }
                const float ax = 1.0f - c.fx, ay = 1.0f - c.fy, az = 1.0f - c.fz;
                // d/dfx: sum over (y,z) corners of wy*wz*(val[x=1]-val[x=0]) etc.
                const float dx = ay * az * (val[1] - val[0]) + c.fy * az * (val[3] - val[2]) +
                                 ay * c.fz * (val[5] - val[4]) + c.fy * c.fz * (val[7] - val[6]);
                const float dy = ax * az * (val[2] - val[0]) + c.fx * az * (val[3] - val[1]) +
                                 ax * c.fz * (val[6] - val[4]) + c.fx * c.fz * (val[7] - val[5]);
                const float dz = ax * ay * (val[4] - val[0]) + c.fx * ay * (val[5] - val[1]) +
                                 ax * c.fy * (val[6] - val[2]) + c.fx * c.fy * (val[7] - val[3]);
                *gx = fmaf(L.scale, dx, *gx);
                *gy = fmaf(L.scale, dy, *gy);
                *gz = fmaf(L.scale, dz, *gz);
            }
            if (grad_table) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float w = corner_weight(c, k);
                    red_add_f32x2(grad_table + 2 * (size_t)idx[k], w * d0, w * d1);
                }
            }
        }
    }
}

// reduce per-point position gradients to their ray: d_o += G/2, d_d += z*G/2
__device__ __forceinline__ void reduce_ray_grad(float* __restrict__ d_rays, bool valid, size_t ray, float z,
                                                float gx, float gy, float gz) {
    float v[6] = {0.5f * gx, 0.5f * gy, 0.5f * gz, 0.5f * z * gx, 0.5f * z * gy, 0.5f * z * gz};
    if (!valid) {
#pragma unroll
        for (int i = 0; i < 6; ++i) v[i] = 0.0f;
    }
    const unsigned long long r0 = __shfl_sync(0xffffffffu, (unsigned long long)ray, 0);
    const bool same = __all_sync(0xffffffffu, !valid || (unsigned long long)ray == r0);
    const bool lane0_valid = __shfl_sync(0xffffffffu, (int)valid, 0) != 0;
    if (same && lane0_valid) {
#pragma unroll
        for (int i = 0; i < 6; ++i) v[i] = warp_sum(v[i]);
        if ((threadIdx.x & 31) == 0) {
#pragma unroll
            for (int i = 0; i < 6; ++i) atomicAdd(d_rays + r0 * LB_RAY_STRIDE + i, v[i]);
        }
    } else if (valid) {
#pragma unroll
        for (int i = 0; i < 6; ++i) atomicAdd(d_rays + ray * LB_RAY_STRIDE + i, v[i]);
    }
}

template <bool kFromRays, bool kInputGrad>
__global__ void __launch_bounds__(kThreads)
sigma_bwd_kernel(lb_level_table T, const float* __restrict__ rays, const float* __restrict__ zv,
                 const float* __restrict__ x01, const float* __restrict__ params,
                 const float* __restrict__ enc_save, const float* __restrict__ d_sigma, size_t n_points,
                 int n_samples, float* __restrict__ grad_params, float* __restrict__ d_rays,
                 float* __restrict__ d_x01) {
    extern __shared__ __align__(16) float sm[];
    float* w1s = sm;                                 // [64][32]
    float* w2s = w1s + LB_MLP_W * LB_MLP_IN;         // [64]
    float* dh_s = w2s + LB_MLP_W;                    // [64][kLd]  dL/dh (after relu mask)
    float* hs_s = dh_s + LB_MLP_W * kLd;             // [64][kLd]  d_sigma * h
    float* e_s = hs_s + LB_MLP_W * kLd;              // [32][kLd]  encodings
    load_mlp_to_smem(params, w1s, w2s);
    __syncthreads();
    const float2* table = reinterpret_cast<const float2*>(params + LB_MLP_PARAMS);
    float* grad_table = grad_params ? grad_params + LB_MLP_PARAMS : nullptr;
    const int tid = threadIdx.x;
    const int j0 = (tid >> 3) * 4, k0 = (tid & 7) * 4;   // this thread's 4x4 block of dW1
    float acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0f;
    float acc_w2 = 0.0f;

    const size_t n_tiles = (n_points + kThreads - 1) / kThreads;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t p = tile * kThreads + tid;
        const bool valid = p < n_points;
        float e[LB_MLP_IN], de[LB_MLP_IN];
        float ds = 0.0f, zval = 0.0f;
        Point q = {0.f, 0.f, 0.f};
        if (valid) {
            q = load_point<kFromRays>(rays, zv, x01, p, n_samples, &zval);
            ds = __ldg(d_sigma + p);
            const float4* src = reinterpret_cast<const float4*>(enc_save + p * LB_MLP_IN);
#pragma unroll
            for (int k4 = 0; k4 < LB_MLP_IN / 4; ++k4) {
                const float4 v = __ldg(src + k4);
                e[4 * k4] = v.x; e[4 * k4 + 1] = v.y; e[4 * k4 + 2] = v.z; e[4 * k4 + 3] = v.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < LB_MLP_IN; ++k) e[k] = 0.0f;
        }
#pragma unroll
        for (int k = 0; k < LB_MLP_IN; ++k) { de[k] = 0.0f; e_s[k * kLd + tid] = e[k]; }
#pragma unroll 2
        for (int j = 0; j < LB_MLP_W; ++j) {
            const float a = hidden_unit(w1s, j, e);
            const float dh = (a > 0.0f) ? ds * w2s[j] : 0.0f;
            hs_s[j * kLd + tid] = ds * fmaxf(a, 0.0f);
            dh_s[j * kLd + tid] = dh;
            const float4* row = reinterpret_cast<const float4*>(w1s + j * LB_MLP_IN);
#pragma unroll
            for (int k4 = 0; k4 < LB_MLP_IN / 4; ++k4) {
                const float4 w = row[k4];
                de[4 * k4 + 0] = fmaf(w.x, dh, de[4 * k4 + 0]);
                de[4 * k4 + 1] = fmaf(w.y, dh, de[4 * k4 + 1]);
                de[4 * k4 + 2] = fmaf(w.z, dh, de[4 * k4 + 2]);
                de[4 * k4 + 3] = fmaf(w.w, dh, de[4 * k4 + 3]);
            }
        }
        float gx = 0.0f, gy = 0.0f, gz = 0.0f;
        if (valid) scatter_point<kInputGrad>(T, table, grad_table, q, de, &gx, &gy, &gz);
        if (kInputGrad) {
            if (kFromRays) {
                reduce_ray_grad(d_rays, valid, valid ? p / (size_t)n_samples : 0, zval, gx, gy, gz);
            } else if (valid) {
                d_x01[3 * p + 0] = gx; d_x01[3 * p + 1] = gy; d_x01[3 * p + 2] = gz;
            }
        }
        __syncthreads();
        if (grad_params) {
            // dW1[j][k] += sum_p dh[p][j] * e[p][k]  (4x4 register block per thread)
            for (int pp = 0; pp < kThreads; ++pp) {
                float a4[4], b4[4];
#pragma unroll
                for (int a = 0; a < 4; ++a) a4[a] = dh_s[(j0 + a) * kLd + pp];
#pragma unroll
                for (int b = 0; b < 4; ++b) b4[b] = e_s[(k0 + b) * kLd + pp];
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(a4[a], b4[b], acc[a][b]);
            }
            if (tid < LB_MLP_W) {
                for (int pp = 0; pp < kThreads; ++pp) acc_w2 += hs_s[tid * kLd + pp];
            }
        }
        __syncthreads();
    }
    if (grad_params) {
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) atomicAdd(grad_params + (j0 + a) * LB_MLP_IN + k0 + b, acc[a][b]);
        if (tid < LB_MLP_W) atomicAdd(grad_params + LB_MLP_W * LB_MLP_IN + tid, acc_w2);
    }
}

__global__ void __launch_bounds__(kThreads)
encode_bwd_kernel(lb_level_table T, const float* __restrict__ x01, const float2* __restrict__ table,
                  const float* __restrict__ d_enc, size_t n_points, float* __restrict__ grad_table,
                  float* __restrict__ d_x01) {
    const int width = T.n_levels * LB_FEATS;
    for (size_t p = (size_t)blockIdx.x * kThreads + threadIdx.x; p < n_points; p += (size_t)gridDim.x * kThreads) {
        const Point q = load_point<false>(nullptr, nullptr, x01, p, 1, nullptr);
        float de[LB_MLP_IN];
#pragma unroll
        for (int k = 0; k < LB_MLP_IN; ++k) de[k] = (k < width) ? __ldg(d_enc + p * width + k) : 0.0f;
        float gx = 0.0f, gy = 0.0f, gz = 0.0f;
        if (d_x01) {
            scatter_point<true>(T, table, grad_table, q, de, &gx, &gy, &gz);
            d_x01[3 * p + 0] = gx; d_x01[3 * p + 1] = gy; d_x01[3 * p + 2] = gz;
        } else {
            scatter_point<false>(T, table, grad_table, q, de, &gx, &gy, &gz);
        }
    }
}

int grid_for(const lb_plan* plan, size_t n_points, int blocks_per_sm) {
    const size_t tiles = (n_points + kThreads - 1) / kThreads;
    const size_t cap = (size_t)plan->sm_count * blocks_per_sm;
    return (int)(tiles < cap ? (tiles ? tiles : 1) : cap);
}

}  // namespace

// ------------------------------------------------------------------------------------------
// C-ABI
// ------------------------------------------------------------------------------------------
extern "C" int lb_encode_fwd(const lb_plan* plan, const float* x01, const float* table, int64_t n_points,
                             float* enc, void* stream) {
    LB_REQUIRE(plan && x01 && table && enc, "null pointer");
    LB_REQUIRE(n_points >= 0, "negative size");
    if (n_points == 0) return LB_OK;
    encode_fwd_kernel<<<grid_for(plan, (size_t)n_points, 8), kThreads, 0, (cudaStream_t)stream>>>(
        lb_make_level_table(plan), x01, reinterpret_cast<const float2*>(table), (size_t)n_points, enc);
    LB_LAUNCH_CHECK("lb_encode_fwd");
    return LB_OK;
}

extern "C" int lb_encode_bwd(const lb_plan* plan, const float* x01, const float* table, const float* d_enc,
                             int64_t n_points, float* grad_table, float* d_x01, void* stream) {
    LB_REQUIRE(plan && x01 && table && d_enc, "null pointer");
    LB_REQUIRE(grad_table || d_x01, "nothing to compute");
    LB_REQUIRE(n_points >= 0, "negative size");
    if (n_points == 0) return LB_OK;
    encode_bwd_kernel<<<grid_for(plan, (size_t)n_points, 8), kThreads, 0, (cudaStream_t)stream>>>(
        lb_make_level_table(plan), x01, reinterpret_cast<const float2*>(table), d_enc, (size_t)n_points,
        grad_table, d_x01);
    LB_LAUNCH_CHECK("lb_encode_bwd");
    return LB_OK;
}

extern "C" int lb_sigma_fwd(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                            const float* params, int n_rays, int n_samples, float* sigma, float* enc_save,
                            void* stream) {
    LB_REQUIRE(plan && params && sigma, "null pointer");
    LB_REQUIRE((rays && z) || x01, "need (rays, z) or x01");
    LB_REQUIRE(plan->cfg.n_levels * LB_FEATS == LB_MLP_IN, "fused sigma head needs 16 levels x 2 features");
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1, "bad sizes");
    const size_t P = (size_t)n_rays * (size_t)n_samples;
    if (P == 0) return LB_OK;
    const lb_level_table T = lb_make_level_table(plan);
    const int grid = grid_for(plan, P, 8);
    if (rays) {
        sigma_fwd_kernel<true><<<grid, kThreads, 0, (cudaStream_t)stream>>>(T, rays, z, nullptr, params, P, n_samples, sigma, enc_save);
    } else {
        sigma_fwd_kernel<false><<<grid, kThreads, 0, (cudaStream_t)stream>>>(T, nullptr, nullptr, x01, params, P, 1, sigma, enc_save);
    }
    LB_LAUNCH_CHECK("lb_sigma_fwd");
    return LB_OK;
}

template <bool A, bool B>
static int launch_sigma_bwd(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                            const float* params, const float* enc_save, const float* d_sigma, size_t P,
                            int n_samples, float* grad_params, float* d_rays, float* d_x01, cudaStream_t st) {
    const size_t smem = (size_t)(LB_MLP_W * LB_MLP_IN + LB_MLP_W + (2 * LB_MLP_W + LB_MLP_IN) * kLd) * sizeof(float);
    int rc = lb_check_cuda(cudaFuncSetAttribute(sigma_bwd_kernel<A, B>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                           "lb_sigma_bwd: smem opt-in");
    if (rc) return rc;
    sigma_bwd_kernel<A, B><<<grid_for(plan, P, 2), kThreads, smem, st>>>(
        lb_make_level_table(plan), rays, z, x01, params, enc_save, d_sigma, P, n_samples, grad_params, d_rays, d_x01);
    LB_LAUNCH_CHECK("lb_sigma_bwd");
    return LB_OK;
}

extern "C" int lb_sigma_bwd(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                            const float* params, const float* enc_save, const float* d_sigma, int n_rays,
                            int n_samples, float* grad_params, float* d_rays, float* d_x01, void* stream) {
    LB_REQUIRE(plan && params && enc_save && d_sigma, "null pointer");
    LB_REQUIRE((rays && z) || x01, "need (rays, z) or x01");
    LB_REQUIRE(plan->cfg.n_levels * LB_FEATS == LB_MLP_IN, "fused sigma head needs 16 levels x 2 features");
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1, "bad sizes");
    LB_REQUIRE(grad_params || d_rays || d_x01, "nothing to compute");
    const size_t P = (size_t)n_rays * (size_t)n_samples;
    if (P == 0) return LB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (rays) {
        LB_REQUIRE(!d_x01, "d_x01 is for the x01 point source");
        if (d_rays) return launch_sigma_bwd<true, true>(plan, rays, z, nullptr, params, enc_save, d_sigma, P, n_samples, grad_params, d_rays, nullptr, st);
        return launch_sigma_bwd<true, false>(plan, rays, z, nullptr, params, enc_save, d_sigma, P, n_samples, grad_params, nullptr, nullptr, st);
    }
    LB_REQUIRE(!d_rays, "d_rays is for the (rays, z) point source");
    if (d_x01) return launch_sigma_bwd<false, true>(plan, nullptr, nullptr, x01, params, enc_save, d_sigma, P, 1, grad_params, nullptr, d_x01, st);
    return launch_sigma_bwd<false, false>(plan, nullptr, nullptr, x01, params, enc_save, d_sigma, P, 1, grad_params, nullptr, nullptr, st);
}
