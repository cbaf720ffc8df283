// R4: the fused sigma head = hash-grid encode (lb_encode.cu) + 32-64-1 MLP (lb_mlp.cu), forward and backward.
//
// Replaces tcnn.NetworkWithInputEncoding(3 -> 1) as used by DecoupledNeRF (reference
// src/models/nerf_tcnn.py:35-38,63-72).  With tables that fit the L2, each direction is ONE warp-specialised kernel for
// either point source -- (rays, z) of the fused render path or x01 [P,3] as tcnn.NetworkWithInputEncoding receives it
// (lb_mlp_umma.cu: gather warps <-> tcgen05 MLP warps through a shared-memory ring); `enc` [P,32] is then written once by
// the forward and read once by the backward.  Otherwise (level-group passes for tables beyond the L2) two launches per
// direction with `enc` as the hand-off, overwritten IN PLACE with dL/d_enc by the MLP backward for the scatter kernel.
#include "lb_common.cuh"

extern "C" int lb_sigma_fwd(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                            const float* params, int n_rays, int n_samples, float* sigma, float* enc,
                            uint64_t* act, void* stream) {
    LB_REQUIRE(plan, "null pointer");
    LB_REQUIRE(n_rays == 0 || (params && sigma), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_rays == 0 || (rays && z) || x01, "need (rays, z) or x01");
    LB_REQUIRE(n_rays == 0 || enc || lb_sigma_fwd_fused_covers(plan),
               "enc [P,32] is required: it is the encode->MLP hand-off unless lb_sigma_fwd_enc_optional(plan)");
    LB_REQUIRE(plan->cfg.n_levels * LB_FEATS == LB_MLP_IN, "fused sigma head needs 16 levels x 2 features");
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1, "bad sizes");
    const size_t P = (size_t)n_rays * (size_t)n_samples;
    LB_REQUIRE(P < (size_t)1 << 32, "n_rays * n_samples must be below 2^32");
    if (P == 0) return LB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (act || !enc) {   // one kernel for both stages, either point source, when the configuration allows it (lb_mlp_umma.cu)
        const int frc = lb_launch_sigma_fwd_fused(plan, rays, z, rays ? nullptr : x01, params, P, n_samples, sigma, enc, act, st);
        if (frc <= 0) return frc;
    }
    // two-kernel path.  Tables beyond the L2 with the tensor-core MLP: `enc` is the level-major hand-off (planes
    // [16][P] float2, same size) written by single-level-group passes (lb_encode.cu); otherwise row-major [P,32].
    if (act && !lb_encode_single_group(plan)) {
        int rc = lb_launch_encode_fwd_lm(plan, rays, z, x01, params + LB_MLP_PARAMS, P, rays ? n_samples : 1, enc, st);
        if (rc) return rc;
        return lb_launch_mlp_fwd(plan, params, enc, P, sigma, act, st, true);
    }
    int rc = lb_launch_encode_fwd(plan, rays, z, x01, params + LB_MLP_PARAMS, P, rays ? n_samples : 1, enc, st);
    if (rc) return rc;
    return lb_launch_mlp_fwd(plan, params, enc, P, sigma, act, st);
}

extern "C" int lb_sigma_fwd_enc_optional(const lb_plan* plan) { return plan && lb_sigma_fwd_fused_covers(plan) ? 1 : 0; }

extern "C" int lb_sigma_bwd(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                            const float* params, float* enc_inout, const uint64_t* act, const float* d_sigma,
                            int n_rays, int n_samples, float* grad_params, float* d_rays, float* d_x01,
                            void* stream) {
    LB_REQUIRE(plan, "null pointer");
    LB_REQUIRE(n_rays == 0 || (params && enc_inout && d_sigma), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_rays == 0 || (rays && z) || x01, "need (rays, z) or x01");
    LB_REQUIRE(plan->cfg.n_levels * LB_FEATS == LB_MLP_IN, "fused sigma head needs 16 levels x 2 features");
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1, "bad sizes");
    LB_REQUIRE(n_rays == 0 || grad_params || d_rays || d_x01, "nothing to compute");
    LB_REQUIRE(!(rays && d_x01) && !(!rays && d_rays), "d_rays goes with (rays, z), d_x01 with x01");
    const size_t P = (size_t)n_rays * (size_t)n_samples;
    LB_REQUIRE(P < (size_t)1 << 32, "n_rays * n_samples must be below 2^32");
    if (P == 0) return LB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    {      // one kernel for both stages, either point source, when the configuration allows it (lb_mlp_umma.cu)
        const int frc = lb_launch_sigma_bwd_fused(plan, rays, z, rays ? nullptr : x01, params, enc_inout, act, d_sigma, P,
                                                  n_samples, grad_params, d_rays, d_x01, st);
        if (frc <= 0) return frc;
    }
    const bool lm = act && !lb_encode_single_group(plan);      // the layout lb_sigma_fwd left in `enc` (see there)
    int rc = lb_launch_mlp_bwd(plan, params, enc_inout, d_sigma, act, P, grad_params, st, lm);
    if (rc) return rc;
    return lb_launch_encode_bwd(plan, rays, z, x01, params + LB_MLP_PARAMS, enc_inout, P, rays ? n_samples : 1,
                                grad_params ? grad_params + LB_MLP_PARAMS : nullptr, d_rays, d_x01, st, lm);
}

// ------------------------------------------------------------------------------------------
// Forward-mode input derivative (pose refinement, reference src/mapping/optimizer.py:701-712 -> 849: the lidar pose is a
// leaf of the graph, so autograd carries dL/dsigma back through tcnn's input gradient to the ray origins / directions).
// lb_sigma_fwd_jvp returns d sigma / d x01 [P,3] with the forward (lb_mlp_umma.cu, sigma_fwd_jvp_fused_kernel), so the
// backward needs no second table gather: lb_sigma_bwd runs without d_rays / d_x01 and lb_rays_grad_from_jvp folds
// dL/dsigma * d sigma / d x01 into the rays.
// ------------------------------------------------------------------------------------------
extern "C" int lb_sigma_fwd_jvp_supported(const lb_plan* plan) { return plan && lb_sigma_fwd_fused_covers(plan) ? 1 : 0; }

extern "C" int lb_sigma_fwd_jvp(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                                const float* params, int n_rays, int n_samples, float* sigma, float* enc,
                                uint64_t* act, float* dsigma_dx, void* stream) {
    LB_REQUIRE(plan, "null pointer");
    LB_REQUIRE(n_rays == 0 || (params && sigma && dsigma_dx), "null pointer");
    LB_REQUIRE(n_rays == 0 || (rays && z) || x01, "need (rays, z) or x01");
    LB_REQUIRE(plan->cfg.n_levels * LB_FEATS == LB_MLP_IN, "fused sigma head needs 16 levels x 2 features");
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1, "bad sizes");
    const size_t P = (size_t)n_rays * (size_t)n_samples;
    LB_REQUIRE(P < (size_t)1 << 32, "n_rays * n_samples must be below 2^32");
    if (P == 0) return LB_OK;
    const int frc = lb_launch_sigma_fwd_jvp_fused(plan, rays, z, rays ? nullptr : x01, params, P, n_samples, sigma, enc, act,
                                                  dsigma_dx, (cudaStream_t)stream);
    if (frc <= 0) return frc;
    lb_set_error("lb_sigma_fwd_jvp: this plan is not covered by the one-kernel forward (lb_sigma_fwd_jvp_supported); use "
                 "lb_sigma_fwd and the d_rays / d_x01 outputs of lb_sigma_bwd");
    return LB_ERR_UNSUPPORTED;
}

namespace {

// one warp per ray: G = sum_s ds[s] g[s], H = sum_s ds[s] z[s] g[s]  ->  d_o += G / 2, d_d += H / 2
// (x01 = (o + d z + 1) / 2; the same 0.5 as sigma_bwd_fused_kernel's d_rays)
__global__ void __launch_bounds__(256) rays_grad_from_jvp_kernel(const float* __restrict__ zv, const float* __restrict__ dsdx,
                                                                 const float* __restrict__ d_sigma, int n_rays, int n_samples,
                                                                 float* __restrict__ d_rays) {
    const int ray = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (ray >= n_rays) return;
    const size_t base = (size_t)ray * n_samples;
    float a[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int s = lane; s < n_samples; s += 32) {
        const float ds = __ldcs(d_sigma + base + s);
        if (ds == 0.0f) continue;                                   // ReLU-inactive or masked samples: most of a young field
        const float zz = __ldcs(zv + base + s);
        const float gx = ds * __ldcs(dsdx + 3 * (base + s)), gy = ds * __ldcs(dsdx + 3 * (base + s) + 1),
                    gz = ds * __ldcs(dsdx + 3 * (base + s) + 2);
        a[0] += gx; a[1] += gy; a[2] += gz;
        a[3] = fmaf(zz, gx, a[3]); a[4] = fmaf(zz, gy, a[4]); a[5] = fmaf(zz, gz, a[5]);
    }
#pragma unroll
    for (int i = 0; i < 6; ++i) a[i] = warp_sum(a[i]);
    if (lane < 6) {
        float v = a[0];
#pragma unroll
        for (int i = 1; i < 6; ++i) v = lane == i ? a[i] : v;
        d_rays[(size_t)ray * LB_RAY_STRIDE + lane] += 0.5f * v;
    }
}

}  // namespace

extern "C" int lb_rays_grad_from_jvp(const float* z, const float* dsigma_dx, const float* d_sigma, int n_rays, int n_samples,
                                     float* d_rays, void* stream) {
    LB_REQUIRE(n_rays == 0 || (z && dsigma_dx && d_sigma && d_rays), "null pointer");
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1, "bad sizes");
    if (n_rays == 0) return LB_OK;
    rays_grad_from_jvp_kernel<<<(n_rays + 7) / 8, 256, 0, (cudaStream_t)stream>>>(z, dsigma_dx, d_sigma, n_rays, n_samples, d_rays);
    return lb_check_cuda(cudaGetLastError(), "rays_grad_from_jvp launch");
}

// ------------------------------------------------------------------------------------------
// lb_sigma_fwd_noalloc: sigma-head inference for callers that cannot allocate device memory (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// lb_sigma_fwd needs two scratch buffers from its caller whenever the plan is not covered by the one-kernel path
// (tables beyond the L2: T >= 2^20): `enc` (the encode -> MLP hand-off, 128 B per point) and `act` (ReLU masks, 8 B per
// point).  Here both live in ONE file-scope device array of fixed size, LB_NOALLOC_MAX_POINTS points = 8.5 MiB, which
// every GPU that loads the library holds until the process exits.  The caller passes inputs and the sigma output only.
//
//   * g_sigma_scratch is a device symbol: in host code its name is a host-side stand-in, NOT a device address, so the
//     launchers are given the pointer cudaGetSymbolAddress returns -- looked up once per device, with that device
//     current, and cached per device ordinal.
//   * nothing in the array has to start at zero: every element a call reads (planes / rows and masks of points
//     < P) is written earlier in the same call; tile tails beyond P are zero-filled by the MLP kernel's own staging.
//     The array keeps the previous call's contents otherwise.
//   * every call on a device shares the array: two calls that overlap on the GPU (different streams, or graphs
//     replayed concurrently) RACE on it.  Serialise them (one stream, or events between streams).
//   * no cudaMalloc / cudaFree / allocating library routine is called here or below (the launchers only launch).
// ------------------------------------------------------------------------------------------
#include <mutex>

namespace {

constexpr size_t kScratchEncBytes = (size_t)LB_NOALLOC_MAX_POINTS * LB_MLP_IN * sizeof(float);     // 8 MiB
constexpr size_t kScratchActBytes = (size_t)LB_NOALLOC_MAX_POINTS * sizeof(uint64_t);               // 0.5 MiB
static __device__ __align__(256) unsigned char g_sigma_scratch[kScratchEncBytes + kScratchActBytes];

constexpr int kMaxDevices = 64;
unsigned char* g_scratch_ptr[kMaxDevices] = {};      // device address of g_sigma_scratch, per device ordinal
std::mutex g_scratch_mutex;

// device address of the scratch array on the CURRENT device (set up on first use per device); nullptr on error
int scratch_for_current_device(unsigned char** out) {
    int dev = -1;
    int rc = lb_check_cuda(cudaGetDevice(&dev), "lb_sigma_fwd_noalloc: cudaGetDevice");
    if (rc) return rc;
    if (dev < 0 || dev >= kMaxDevices) {
        lb_set_error("lb_sigma_fwd_noalloc: device ordinal %d outside [0, %d)", dev, kMaxDevices);
        return LB_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        rc = lb_check_cuda(cudaGetSymbolAddress(&p, g_sigma_scratch), "lb_sigma_fwd_noalloc: cudaGetSymbolAddress(scratch)");
        if (rc) return rc;
        if (!p) {
            lb_set_error("lb_sigma_fwd_noalloc: cudaGetSymbolAddress returned a null device address");
            return LB_ERR_CUDA;
        }
        g_scratch_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_ptr[dev];
    return LB_OK;
}

}  // namespace

extern "C" int lb_sigma_fwd_noalloc(const lb_plan* plan, const float* rays, const float* z, const float* x01,
                                    const float* params, int n_rays, int n_samples, float* sigma, void* stream) {
    LB_REQUIRE(plan, "null pointer");
    LB_REQUIRE(n_rays == 0 || (params && sigma), "null pointer");
    LB_REQUIRE(n_rays == 0 || (rays && z) || x01, "need (rays, z) or x01");
    LB_REQUIRE(plan->cfg.n_levels * LB_FEATS == LB_MLP_IN, "fused sigma head needs 16 levels x 2 features");
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1, "bad sizes");
    const size_t P = (size_t)n_rays * (size_t)(rays ? n_samples : 1);
    if (P > (size_t)LB_NOALLOC_MAX_POINTS) {
        lb_set_error("lb_sigma_fwd_noalloc: %zu sample points exceed the built-in scratch of LB_NOALLOC_MAX_POINTS = %d "
                     "points; split the batch or call lb_sigma_fwd with caller-owned enc / act buffers",
                     P, (int)LB_NOALLOC_MAX_POINTS);
        return LB_ERR_BAD_ARGUMENT;
    }
    if (P == 0) return LB_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int ns = rays ? n_samples : 1;
    // plans the one-kernel path covers need no hand-off at all in inference: gather + MLP in one launch, nothing saved
    {
        const int frc = lb_launch_sigma_fwd_fused(plan, rays, z, rays ? nullptr : x01, params, P, ns, sigma, nullptr, nullptr, st);
        if (frc <= 0) return frc;
    }
    unsigned char* scratch = nullptr;
    int rc = scratch_for_current_device(&scratch);
    if (rc) return rc;
    float* enc = reinterpret_cast<float*>(scratch);
    uint64_t* act = reinterpret_cast<uint64_t*>(scratch + kScratchEncBytes);
    // two-kernel path through the level-major hand-off (planes [16][P] float2), exactly as lb_sigma_fwd runs it
    rc = lb_launch_encode_fwd_lm(plan, rays, z, x01, params + LB_MLP_PARAMS, P, ns, enc, st);
    if (rc) return rc;
    return lb_launch_mlp_fwd(plan, params, enc, P, sigma, act, st, true);
}
