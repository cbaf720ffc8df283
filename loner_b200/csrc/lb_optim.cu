// O1/O2: fused Adam / SGD steps and the occupancy-grid log-odds update.
//
// Replaces torch.optim.Adam over the flat tcnn parameter vector (reference
// src/mapping/optimizer.py:255-265,460-464), and _step_occupancy_grid (optimizer.py:897-908) with
// get_logits_grad (src/models/losses.py:54-62) and the backward of F.grid_sample
// (src/models/model_tcnn.py:126-134).
#include "lb_common.cuh"

namespace {

__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, size_t n,
            float lr_over_bc1, float inv_sqrt_bc2, float b1, float b2, float omb1, float omb2, float eps,
            float gscale, int zero_grad) {
    const size_t n4 = n / 4;
    float4* p4 = reinterpret_cast<float4*>(p);
    float4* g4 = reinterpret_cast<float4*>(g);
    float4* m4 = reinterpret_cast<float4*>(m);
    float4* v4 = reinterpret_cast<float4*>(v);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    auto upd = [&](float& pp, float gg, float& mm, float& vv) {
        gg *= gscale;
        mm = b1 * mm + omb1 * gg;
        vv = b2 * vv + omb2 * gg * gg;
        const float denom = sqrtf(vv) * inv_sqrt_bc2 + eps;
        pp -= lr_over_bc1 * (mm / denom);
    };
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 P = p4[i], G = g4[i], M = m4[i], V = v4[i];
        upd(P.x, G.x, M.x, V.x); upd(P.y, G.y, M.y, V.y); upd(P.z, G.z, M.z, V.z); upd(P.w, G.w, M.w, V.w);
        p4[i] = P; m4[i] = M; v4[i] = V;
        if (zero_grad) g4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        float P = p[i], M = m[i], V = v[i];
        upd(P, g[i], M, V);
        p[i] = P; m[i] = M; v[i] = V;
        if (zero_grad) g[i] = 0.0f;
    }
}

__global__ void __launch_bounds__(256)
sgd_kernel(float* __restrict__ p, float* __restrict__ g, size_t n, float lr, int zero_grad) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        p[i] -= lr * g[i];
        if (zero_grad) g[i] = 0.0f;
    }
}

// occ_grad[corner] += w_corner * u, u = +0.25 if zm-g < -2, -2.5 if |zm-g| < 2 (strict), else 0
//
// A warp takes 32 consecutive samples.  Samples are depth-sorted along their ray and a 100^3 voxel holds several of
// them (the sampler concentrates them at the surfaces), so neighbouring lanes mostly address the same 8 corners: runs
// of lanes with the same base cell are summed by a segmented shuffle reduction and only the run head issues the
// atomics (the grid is 4 MB, the cost of this kernel is same-address L2 atomics).
__global__ void __launch_bounds__(256)
occ_grad_kernel(const float* __restrict__ rays, const float* __restrict__ zv, const float* __restrict__ depth_gt,
                float scale, size_t n_points, int S, int D, float* __restrict__ occ_grad) {
    const unsigned kFull = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    const size_t n_round = (n_points + 31) & ~(size_t)31;          // whole warps stay in the loop together
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_round; p += stride) {
        float u = 0.0f, z = 0.0f;
        const float* ray = rays;
        if (p < n_points) {
            const size_t r = p / (size_t)S;
            ray = rays + r * LB_RAY_STRIDE;
            z = __ldg(zv + p);
            const float x = z * scale - __ldg(depth_gt + r) * scale;
            if (x < -2.0f) u = 0.25f;
            else if (x > -2.0f && x < 2.0f) u = -2.5f;
        }
        const bool active = (u != 0.0f);
        if (!__any_sync(kFull, active)) continue;
        const float px = __ldg(ray + 0) + __ldg(ray + 3) * z;
        const float py = __ldg(ray + 1) + __ldg(ray + 4) * z;
        const float pz = __ldg(ray + 2) + __ldg(ray + 5) * z;
        const float fD = (float)D;
        const float ix = ((px + 1.0f) * fD - 1.0f) * 0.5f, iy = ((py + 1.0f) * fD - 1.0f) * 0.5f,
                    iz = ((pz + 1.0f) * fD - 1.0f) * 0.5f;
        const float x0f = floorf(ix), y0f = floorf(iy), z0f = floorf(iz);
        const float fx = ix - x0f, fy = iy - y0f, fz = iz - z0f;
        // clamp far-out cells to one sentinel layer outside the grid: no corner of such a cell is inside anyway
        const int x0 = (int)fminf(fmaxf(x0f, -2.0f), (float)D), y0 = (int)fminf(fmaxf(y0f, -2.0f), (float)D),
                  z0 = (int)fminf(fmaxf(z0f, -2.0f), (float)D);
        float w[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const int bx = c & 1, by = (c >> 1) & 1, bz = (c >> 2) & 1;
            w[c] = active ? (bx ? fx : 1.0f - fx) * (by ? fy : 1.0f - fy) * (bz ? fz : 1.0f - fz) * u : 0.0f;
        }
        // runs of equal base cells among consecutive lanes (inactive lanes join their neighbour's run with zeros)
        const int px0 = __shfl_up_sync(kFull, x0, 1), py0 = __shfl_up_sync(kFull, y0, 1), pz0 = __shfl_up_sync(kFull, z0, 1);
        const bool same = lane > 0 && px0 == x0 && py0 == y0 && pz0 == z0;
        const unsigned heads = __ballot_sync(kFull, !same);
        if (heads != kFull) {
            const unsigned rest = (lane < 31) ? (heads >> (lane + 1)) : 0u;
            const int dist = rest ? __ffs(rest) : 32 - lane;          // lanes until the next head
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const float t = __shfl_down_sync(kFull, w[c], off);
                    if (off < dist) w[c] += t;
                }
            }
        }
        if (same) continue;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
            const int bx = c & 1, by = (c >> 1) & 1, bz = (c >> 2) & 1;
            const int X = x0 + bx, Y = y0 + by, Z = z0 + bz;
            if (w[c] != 0.0f && X >= 0 && X < D && Y >= 0 && Y < D && Z >= 0 && Z < D)
                atomicAdd(occ_grad + ((size_t)Z * D + Y) * D + X, w[c]);
        }
    }
}

int blocks_for(size_t n, int per_block) {
    size_t b = (n + per_block - 1) / per_block;
    const size_t cap = 148 * 16;
    return (int)(b < cap ? (b ? b : 1) : cap);
}

}  // namespace

extern "C" int lb_adam_step(float* p, float* g, float* m, float* v, int64_t n, double lr, double beta1,
                            double beta2, double eps, int step, float grad_scale, int zero_grad, void* stream) {
    LB_REQUIRE(n == 0 || (p && g && m && v), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n >= 0 && step >= 1, "bad size / step");
    LB_REQUIRE((((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0, "buffers must be 16-byte aligned");
    if (n == 0) return LB_OK;
    // hyper-parameters arrive as doubles (Python floats) so that 1-beta, lr/bc1, sqrt(bc2) round once,
    // exactly like torch.optim.Adam's host-side scalar arithmetic
    const double bc1 = 1.0 - pow(beta1, (double)step);
    const double bc2 = 1.0 - pow(beta2, (double)step);
    adam_kernel<<<blocks_for((size_t)n / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>(
        p, g, m, v, (size_t)n, (float)(lr / bc1), (float)(1.0 / sqrt(bc2)), (float)beta1, (float)beta2,
        (float)(1.0 - beta1), (float)(1.0 - beta2), (float)eps, grad_scale, zero_grad);
    LB_LAUNCH_CHECK("lb_adam_step");
    return LB_OK;
}

extern "C" int lb_sgd_step(float* p, float* g, int64_t n, float lr, int zero_grad, void* stream) {
    LB_REQUIRE(n == 0 || (p && g), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n >= 0, "negative size");
    if (n == 0) return LB_OK;
    sgd_kernel<<<blocks_for((size_t)n, 256), 256, 0, (cudaStream_t)stream>>>(p, g, (size_t)n, lr, zero_grad);
    LB_LAUNCH_CHECK("lb_sgd_step");
    return LB_OK;
}

extern "C" int lb_occ_grad(const float* rays, const float* z, const float* depth_gt, float scale_factor,
                           int n_rays, int n_samples, int occ_dim, float* occ_grad, void* stream) {
    LB_REQUIRE(n_rays == 0 || (rays && z && depth_gt && occ_grad), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1 && occ_dim >= 1, "bad sizes");
    const size_t P = (size_t)n_rays * (size_t)n_samples;
    if (P == 0) return LB_OK;
    occ_grad_kernel<<<blocks_for(P, 256), 256, 0, (cudaStream_t)stream>>>(rays, z, depth_gt, scale_factor, P, n_samples,
                                                                         occ_dim, occ_grad);
    LB_LAUNCH_CHECK("lb_occ_grad");
    return LB_OK;
}
