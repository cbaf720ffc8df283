// Shared device/host helpers for libloner_b200 (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/loner_b200.h"

#define LB_FEATS 2               // features per level (float2 entries)
#define LB_MLP_IN 32             // n_levels * LB_FEATS for the sigma head
#define LB_MLP_W 64              // hidden width
#define LB_MLP_OUT_PAD 16        // tcnn pads the output layer to 16 rows; row 0 is sigma
#define LB_MLP_PARAMS (LB_MLP_W * LB_MLP_IN + LB_MLP_OUT_PAD * LB_MLP_W)

struct lb_level {
    float    scale;      // 2^(l*log2(pls)) * base - 1
    uint32_t res;        // ceil(scale) + 1
    uint32_t res2;       // res * res (z stride of a dense level)
    uint32_t offset;     // in entries (float2)
    uint32_t size;       // entries in this level
    uint32_t dense;      // 1: x + y*res + z*res^2 ; 0: coherent prime hash
    uint32_t mask;       // size-1 when size is a power of two (index & mask), else 0 (index % size)
};

struct lb_level_table {
    lb_level lv[16];
    int32_t  n_levels;
};

struct lb_plan {
    lb_grid_config cfg;
    lb_level       levels[LB_MAX_LEVELS];
    int64_t        n_table_entries;   // float2 entries over all levels
    int64_t        n_mlp_params;
    int            sm_count;
    int64_t        l2_bytes;          // device L2 size (level-group passes for tables that do not fit, lb_encode.cu)
};

void lb_set_error(const char* fmt, ...);
int  lb_check_cuda(cudaError_t e, const char* what);

#define LB_REQUIRE(cond, msg)                                     \
    do {                                                          \
        if (!(cond)) {                                            \
            lb_set_error("%s: %s", __func__, msg);                \
            return LB_ERR_BAD_ARGUMENT;                           \
        }                                                         \
    } while (0)

#define LB_LAUNCH_CHECK(what)                                     \
    do {                                                          \
        int _rc = lb_check_cuda(cudaGetLastError(), what);        \
        if (_rc) return _rc;                                      \
    } while (0)

// internal launchers (lb_encode.cu, lb_mlp.cu) used by the fused entry points in lb_field.cu
int lb_launch_encode_fwd(const lb_plan* plan, const float* rays, const float* z, const float* x01, const float* table,
                         size_t n_points, int n_samples, float* enc, cudaStream_t st);
// level_major: d_enc holds planes [n_levels][n_points] (float2) instead of rows [n_points][32] (lb_encode.cu)
int lb_launch_encode_bwd(const lb_plan* plan, const float* rays, const float* z, const float* x01, const float* table,
                         const float* d_enc, size_t n_points, int n_samples, float* grad_table, float* d_rays,
                         float* d_x01, cudaStream_t st, bool level_major);
int lb_launch_encode_fwd_lm(const lb_plan* plan, const float* rays, const float* z, const float* x01, const float* table,
                            size_t n_points, int n_samples, float* planes, cudaStream_t st);
// act [P] uint64 = ReLU bit masks (hand-off forward -> backward of the tensor-core path); NULL selects the
// fp32 CUDA-core kernels
// level_major: enc is stored as planes [16][n_points] (float2), the hand-off layout of multi-group plans
int lb_launch_mlp_fwd(const lb_plan* plan, const float* mlp, const float* enc, size_t n_points, float* sigma,
                      uint64_t* act, cudaStream_t st, bool level_major = false);
int lb_launch_mlp_bwd(const lb_plan* plan, const float* mlp, float* enc_inout, const float* d_sigma,
                      const uint64_t* act, size_t n_points, float* grad_mlp, cudaStream_t st, bool level_major = false);
int lb_launch_mlp_fwd_umma(const lb_plan* plan, const float* mlp, const float* enc, size_t n_points, float* sigma,
                           uint64_t* act, cudaStream_t st, bool level_major = false);
int lb_launch_mlp_bwd_umma(const lb_plan* plan, const float* mlp, float* enc_inout, const float* d_sigma,
                           const uint64_t* act, size_t n_points, float* grad_mlp, cudaStream_t st, bool level_major = false);

// lb_encode.cu: leading dense levels of a "dense prefix + power-of-two hashed rest" plan (else -1); whether all levels
// fit one L2-resident group
int  lb_encode_dense_prefix(const lb_plan* plan);
bool lb_encode_single_group(const lb_plan* plan);
// lb_mlp_umma.cu: MLP backward + encode backward in one kernel (dL/denc never leaves the SM).  Returns 1 when the
// configuration is not covered (caller runs the two separate kernels), 0 on success, < 0 on error.
bool lb_sigma_fwd_fused_covers(const lb_plan* plan);
int lb_launch_sigma_fwd_fused(const lb_plan* plan, const float* rays, const float* z, const float* x01, const float* params,
                              size_t n_points, int n_samples, float* sigma, float* enc, uint64_t* act, cudaStream_t st);
// forward + d sigma / d x01 [P,3] (forward-mode input derivative); same return convention
int lb_launch_sigma_fwd_jvp_fused(const lb_plan* plan, const float* rays, const float* z, const float* x01, const float* params,
                                  size_t n_points, int n_samples, float* sigma, float* enc, uint64_t* act, float* dsigma_dx,
                                  cudaStream_t st);
int lb_launch_sigma_bwd_fused(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                              const float* params, const float* enc, const uint64_t* act, const float* d_sigma,
                              size_t n_points, int n_samples, float* grad_params, float* d_rays, float* d_x01,
                              cudaStream_t st);

static inline lb_level_table lb_make_level_table(const lb_plan* p) {
    lb_level_table t;
    memset(&t, 0, sizeof(t));
    t.n_levels = p->cfg.n_levels;
    for (int i = 0; i < p->cfg.n_levels && i < 16; ++i) t.lv[i] = p->levels[i];
    return t;
}

// ------------------------------------------------------------------------------------------
// device helpers
// ------------------------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t lb_grid_index(const lb_level& L, uint32_t x, uint32_t y, uint32_t z) {
    uint32_t idx;
    if (L.dense) {
        idx = x + y * L.res + z * L.res * L.res;
    } else {
        idx = x ^ (y * 2654435761u) ^ (z * 805459861u);
    }
    return L.mask ? (idx & L.mask) : (idx % L.size);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// vector reduction (no return value) into global memory: one 8-byte L2 atomic for a float2 entry
__device__ __forceinline__ void red_add_f32x2(float* addr, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(addr), "f"(a), "f"(b) : "memory");
}

#endif  // __CUDACC__
