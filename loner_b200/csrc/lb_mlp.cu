// R4 (network half): the 32 -> 64 -> 1 sigma MLP (no bias, ReLU hidden, linear output), forward and
// backward, as stand-alone streaming kernels over the encodings [P,32].
//
// Replaces tcnn FullyFusedMLP inside tcnn.NetworkWithInputEncoding (reference src/models/nerf_tcnn.py:35-38,
// cfg/nerf_config/default_nerf_hash.yaml:26-31).  Weight layout = tcnn's: W1 [64,32] then W2 [16,64]
// row-major [out,in]; only W2 row 0 produces sigma, rows 1-15 are padding and receive zero gradient.
//
// The shipped library has ONE implementation: the tcgen05 tensor-core kernels of lb_mlp_umma.cu (the caller supplies the
// `act` ReLU-mask buffer).  The fp32 CUDA-core kernels in this file are a cross-check for the tests and exist only in a
// build with -DLB_TEST_SWITCHES (`python -m loner_b200.build --test-switches`; lb_build_flags() & 1): there, act == NULL
// or LONER_B200_MLP=simt selects them.
#include <stdlib.h>

#include "lb_common.cuh"

#ifdef LB_TEST_SWITCHES
namespace {

constexpr int kThreads = 128;
constexpr int kLd = kThreads + 1;   // padded leading dimension of [row][point] staging tiles

__device__ __forceinline__ void load_mlp_to_smem(const float* __restrict__ mlp, float* w1s, float* w2s) {
    for (int i = threadIdx.x; i < LB_MLP_W * LB_MLP_IN; i += blockDim.x) w1s[i] = __ldg(mlp + i);
    for (int i = threadIdx.x; i < LB_MLP_W; i += blockDim.x) w2s[i] = __ldg(mlp + LB_MLP_W * LB_MLP_IN + i);
}

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

__device__ __forceinline__ void load_row(const float* __restrict__ enc, size_t p, float* e) {
    const float4* src = reinterpret_cast<const float4*>(enc + p * LB_MLP_IN);
#pragma unroll
    for (int k4 = 0; k4 < LB_MLP_IN / 4; ++k4) {
        const float4 v = __ldg(src + k4);
        e[4 * k4] = v.x; e[4 * k4 + 1] = v.y; e[4 * k4 + 2] = v.z; e[4 * k4 + 3] = v.w;
    }
}

__global__ void __launch_bounds__(kThreads)
mlp_fwd_simt_kernel(const float* __restrict__ mlp, const float* __restrict__ enc, size_t n_points,
                    float* __restrict__ sigma) {
    __shared__ __align__(16) float w1s[LB_MLP_W * LB_MLP_IN];
    __shared__ float w2s[LB_MLP_W];
    load_mlp_to_smem(mlp, w1s, w2s);
    __syncthreads();
    for (size_t p = (size_t)blockIdx.x * kThreads + threadIdx.x; p < n_points; p += (size_t)gridDim.x * kThreads) {
        float e[LB_MLP_IN];
        load_row(enc, p, e);
        float s = 0.0f;
#pragma unroll 4
        for (int j = 0; j < LB_MLP_W; ++j) s = fmaf(w2s[j], fmaxf(hidden_unit(w1s, j, e), 0.0f), s);
        sigma[p] = s;
    }
}

// d_enc (written IN PLACE over enc) = W1^T (d_sigma * w2 * [h > 0]);  dW1 += dh (x) e;  dW2[0] += d_sigma * relu(h)
__global__ void __launch_bounds__(kThreads)
mlp_bwd_simt_kernel(const float* __restrict__ mlp, float* __restrict__ enc, const float* __restrict__ d_sigma,
                    size_t n_points, float* __restrict__ grad_mlp) {
    extern __shared__ __align__(16) float sm[];
    float* w1s = sm;                                 // [64][32]
    float* w2s = w1s + LB_MLP_W * LB_MLP_IN;         // [64]
    float* dh_s = w2s + LB_MLP_W;                    // [64][kLd]
    float* hs_s = dh_s + LB_MLP_W * kLd;             // [64][kLd]
    float* e_s = hs_s + LB_MLP_W * kLd;              // [32][kLd]
    load_mlp_to_smem(mlp, w1s, w2s);
    __syncthreads();
    const int tid = threadIdx.x;
    const int j0 = (tid >> 3) * 4, k0 = (tid & 7) * 4;
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
        float ds = 0.0f;
        if (valid) {
            ds = __ldg(d_sigma + p);
            load_row(enc, p, e);
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
        if (valid) {
            float4* dst = reinterpret_cast<float4*>(enc + p * LB_MLP_IN);
#pragma unroll
            for (int k4 = 0; k4 < LB_MLP_IN / 4; ++k4) dst[k4] = make_float4(de[4 * k4], de[4 * k4 + 1], de[4 * k4 + 2], de[4 * k4 + 3]);
        }
        __syncthreads();
        if (grad_mlp) {
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
    if (grad_mlp) {
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) atomicAdd(grad_mlp + (j0 + a) * LB_MLP_IN + k0 + b, acc[a][b]);
        if (tid < LB_MLP_W) atomicAdd(grad_mlp + LB_MLP_W * LB_MLP_IN + tid, acc_w2);
    }
}

int grid_for(int sm_count, size_t n_points, int blocks_per_sm) {
    const size_t tiles = (n_points + kThreads - 1) / kThreads;
    const size_t cap = (size_t)sm_count * blocks_per_sm;
    return (int)(tiles < cap ? (tiles ? tiles : 1) : cap);
}

}  // namespace

// LONER_B200_MLP=simt forces the fp32 CUDA-core kernels (test builds only)
static bool use_simt() {
    const char* e = getenv("LONER_B200_MLP");
    return e && strcmp(e, "simt") == 0;
}

int lb_launch_mlp_fwd(const lb_plan* plan, const float* mlp, const float* enc, size_t n_points, float* sigma,
                      uint64_t* act, cudaStream_t st, bool level_major) {
    if (act && (!use_simt() || level_major)) return lb_launch_mlp_fwd_umma(plan, mlp, enc, n_points, sigma, act, st, level_major);
    if (level_major) { lb_set_error("the fp32 CUDA-core MLP reads row-major encodings only"); return LB_ERR_UNSUPPORTED; }
    mlp_fwd_simt_kernel<<<grid_for(plan->sm_count, n_points, 8), kThreads, 0, st>>>(mlp, enc, n_points, sigma);
    return lb_check_cuda(cudaGetLastError(), "mlp_fwd launch");
}

int lb_launch_mlp_bwd(const lb_plan* plan, const float* mlp, float* enc_inout, const float* d_sigma,
                      const uint64_t* act, size_t n_points, float* grad_mlp, cudaStream_t st, bool level_major) {
    if (act && (!use_simt() || level_major)) return lb_launch_mlp_bwd_umma(plan, mlp, enc_inout, d_sigma, act, n_points, grad_mlp, st, level_major);
    if (level_major) { lb_set_error("the fp32 CUDA-core MLP reads row-major encodings only"); return LB_ERR_UNSUPPORTED; }
    const size_t smem = (size_t)(LB_MLP_W * LB_MLP_IN + LB_MLP_W + (2 * LB_MLP_W + LB_MLP_IN) * kLd) * sizeof(float);
    int rc = lb_check_cuda(cudaFuncSetAttribute(mlp_bwd_simt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                           "mlp_bwd smem opt-in");
    if (rc) return rc;
    mlp_bwd_simt_kernel<<<grid_for(plan->sm_count, n_points, 2), kThreads, smem, st>>>(mlp, enc_inout, d_sigma, n_points, grad_mlp);
    return lb_check_cuda(cudaGetLastError(), "mlp_bwd launch");
}
#else   // shipped build: tensor cores only
int lb_launch_mlp_fwd(const lb_plan* plan, const float* mlp, const float* enc, size_t n_points, float* sigma,
                      uint64_t* act, cudaStream_t st, bool level_major) {
    if (!act) {
        lb_set_error("lb_sigma_fwd: the act [P] ReLU-mask buffer is required (the fp32 CUDA-core MLP is a test-only build option)");
        return LB_ERR_UNSUPPORTED;
    }
    return lb_launch_mlp_fwd_umma(plan, mlp, enc, n_points, sigma, act, st, level_major);
}

int lb_launch_mlp_bwd(const lb_plan* plan, const float* mlp, float* enc_inout, const float* d_sigma,
                      const uint64_t* act, size_t n_points, float* grad_mlp, cudaStream_t st, bool level_major) {
    if (!act) {
        lb_set_error("lb_sigma_bwd: the act [P] ReLU-mask buffer is required (the fp32 CUDA-core MLP is a test-only build option)");
        return LB_ERR_UNSUPPORTED;
    }
    return lb_launch_mlp_bwd_umma(plan, mlp, enc_inout, d_sigma, act, n_points, grad_mlp, st, level_major);
}
#endif
