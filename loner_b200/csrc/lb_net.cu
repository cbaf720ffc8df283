// RGB branch, remaining pieces: spherical-harmonics direction encoding (forward / input gradient), colour
// compositing, and the C entry point of the fused network's backward (the tcgen05 kernels are in lb_net_umma.cu).
//
// Replaces tcnn.Encoding(3, {"otype": "SphericalHarmonics", "degree": 4}) (reference src/models/nerf_tcnn.py:43,86),
// the autograd backward of tcnn.Network (nerf_tcnn.py:50-52,89-91) and the colour lines of raw2outputs
// (src/models/rendering_tcnn.py:284-289).
#include "lb_common.cuh"

int lb_launch_net_bwd_umma(const float* x, size_t n_points, int n_in, int n_hidden, int n_out, int out_act, const float* params,
                           const float* acts, const float* out, const float* d_out, float* grad_params, float* d_x,
                           cudaStream_t st);

namespace {

// ------------------------------------------------------------------------------------------
// SH degree 4 on v = 2 * d01 - 1 (16 coefficients; tcnn's real SH basis and ordering)
// ------------------------------------------------------------------------------------------
constexpr float kA = 0.48860251190291987f, kB = 1.0925484305920792f, kC = 0.94617469575755997f,
                kD = 0.31539156525251999f, kE = 0.54627421529603959f, kF = 0.59004358992664352f,
                kG = 2.8906114426405538f, kH = 0.45704579946446572f, kI = 0.3731763325901154f,
                kJ = 1.4453057213202769f;

__global__ void __launch_bounds__(256)
sh4_fwd_kernel(const float* __restrict__ d01, size_t n, float* __restrict__ out) {
    const size_t p = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (p >= n) return;
    const float x = 2.0f * d01[3 * p] - 1.0f, y = 2.0f * d01[3 * p + 1] - 1.0f, z = 2.0f * d01[3 * p + 2] - 1.0f;
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    float4* o = reinterpret_cast<float4*>(out + 16 * p);
    o[0] = make_float4(0.28209479177387814f, -kA * y, kA * z, -kA * x);
    o[1] = make_float4(kB * xy, -kB * yz, kC * z2 - kD, -kB * xz);
    o[2] = make_float4(kE * x2 - kE * y2, kF * y * (-3.0f * x2 + y2), kG * xy * z, kH * y * (1.0f - 5.0f * z2));
    o[3] = make_float4(kI * z * (5.0f * z2 - 3.0f), kH * x * (1.0f - 5.0f * z2), kJ * z * (x2 - y2),
                       kF * x * (-x2 + 3.0f * y2));
}

__global__ void __launch_bounds__(256)
sh4_bwd_kernel(const float* __restrict__ d01, const float* __restrict__ d_out, size_t n, float* __restrict__ d_d01) {
    const size_t p = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (p >= n) return;
    const float x = 2.0f * d01[3 * p] - 1.0f, y = 2.0f * d01[3 * p + 1] - 1.0f, z = 2.0f * d01[3 * p + 2] - 1.0f;
    const float x2 = x * x, y2 = y * y, z2 = z * z;
    const float4* gp = reinterpret_cast<const float4*>(d_out + 16 * p);
    const float4 g0 = gp[0], g1 = gp[1], g2 = gp[2], g3 = gp[3];
    const float q = 1.0f - 5.0f * z2, t = 3.0f * (y2 - x2);
    float gx = -kA * g0.w + kB * y * g1.x - kB * z * g1.w + 2.0f * kE * x * g2.x - 6.0f * kF * x * y * g2.y +
               kG * y * z * g2.z + kH * q * g3.y + 2.0f * kJ * z * x * g3.z + kF * t * g3.w;
    float gy = -kA * g0.y + kB * x * g1.x - kB * z * g1.y - 2.0f * kE * y * g2.x + kF * t * g2.y + kG * x * z * g2.z +
               kH * q * g2.w - 2.0f * kJ * z * y * g3.z + 6.0f * kF * x * y * g3.w;
    float gz = kA * g0.z - kB * y * g1.y + 2.0f * kC * z * g1.z - kB * x * g1.w + kG * x * y * g2.z -
               10.0f * kH * y * z * g2.w + kI * (15.0f * z2 - 3.0f) * g3.x - 10.0f * kH * x * z * g3.y +
               kJ * (x2 - y2) * g3.z;
    d_d01[3 * p] = 2.0f * gx;
    d_d01[3 * p + 1] = 2.0f * gy;
    d_d01[3 * p + 2] = 2.0f * gz;
}

// ------------------------------------------------------------------------------------------
// colour compositing: rgb[n, c] = sum_s w[n, s] col[n, s, c] (+ 1 - sum_s w[n, s] on a white background)
// one warp per ray
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
composite_rgb_fwd_kernel(const float* __restrict__ w, const float* __restrict__ col, int n_rays, int n_samples, int nc,
                         int white_bkgd, float* __restrict__ rgb) {
    const int ray = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (ray >= n_rays) return;
    float acc[4] = {0.f, 0.f, 0.f, 0.f}, ws = 0.0f;
    for (int s = lane; s < n_samples; s += 32) {
        const float wv = __ldg(w + (size_t)ray * n_samples + s);
        ws += wv;
        for (int c = 0; c < nc; ++c) acc[c] = fmaf(wv, __ldg(col + ((size_t)ray * n_samples + s) * nc + c), acc[c]);
    }
    ws = warp_sum(ws);
    for (int c = 0; c < nc; ++c) {
        const float a = warp_sum(acc[c]);
        if (lane == 0) rgb[(size_t)ray * nc + c] = white_bkgd ? a + 1.0f - ws : a;
    }
}

__global__ void __launch_bounds__(256)
composite_rgb_bwd_kernel(const float* __restrict__ w, const float* __restrict__ col, const float* __restrict__ g_rgb,
                         int n_rays, int n_samples, int nc, int white_bkgd, float* __restrict__ d_w,
                         float* __restrict__ d_col) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= (size_t)n_rays * n_samples) return;
    const size_t ray = i / n_samples;
    const float wv = __ldg(w + i);
    float dw = 0.0f;
    for (int c = 0; c < nc; ++c) {
        const float g = __ldg(g_rgb + ray * nc + c);
        dw = fmaf(g, __ldg(col + i * nc + c) - (white_bkgd ? 1.0f : 0.0f), dw);
        if (d_col) d_col[i * nc + c] = g * wv;
    }
    if (d_w) d_w[i] = dw;
}

}  // namespace

extern "C" int lb_sh4_fwd(const float* d01, int64_t n_points, float* out, void* stream) {
    LB_REQUIRE(n_points == 0 || (d01 && out), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_points >= 0, "negative size");
    if (n_points == 0) return LB_OK;
    sh4_fwd_kernel<<<(unsigned)((n_points + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d01, (size_t)n_points, out);
    LB_LAUNCH_CHECK("lb_sh4_fwd");
    return LB_OK;
}

extern "C" int lb_sh4_bwd(const float* d01, const float* d_out, int64_t n_points, float* d_d01, void* stream) {
    LB_REQUIRE(n_points == 0 || (d01 && d_out && d_d01), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_points >= 0, "negative size");
    if (n_points == 0) return LB_OK;
    sh4_bwd_kernel<<<(unsigned)((n_points + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d01, d_out, (size_t)n_points,
                                                                                        d_d01);
    LB_LAUNCH_CHECK("lb_sh4_bwd");
    return LB_OK;
}

extern "C" int lb_net_bwd(const float* x, int64_t n_points, int n_in, int n_hidden, int n_out, int out_act,
                          const float* params, const float* acts, const float* out, const float* d_out,
                          float* grad_params, float* d_x, void* stream) {
    LB_REQUIRE(n_points == 0 || (x && params && acts && d_out), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(grad_params || d_x, "nothing to compute");
    LB_REQUIRE(n_points >= 0, "negative size");
    LB_REQUIRE(n_in >= 16 && n_in <= 64 && n_in % 16 == 0, "n_in must be 16, 32, 48 or 64");
    LB_REQUIRE(n_hidden >= 1 && n_hidden <= 6, "n_hidden_layers must be in [1, 6]");
    LB_REQUIRE(n_out >= 1 && n_out <= 16, "n_out must be in [1, 16]");
    LB_REQUIRE(out_act == 0 || (out_act == 1 && out), "out_act 1 (sigmoid) needs the forward output");
    if (n_points == 0) return LB_OK;
    const int rc = lb_launch_net_bwd_umma(x, (size_t)n_points, n_in, n_hidden, n_out, out_act, params, acts, out, d_out,
                                          grad_params, d_x, (cudaStream_t)stream);
    if (rc > 0) {
        lb_set_error("lb_net_bwd: configuration does not fit in shared memory");
        return LB_ERR_UNSUPPORTED;
    }
    return rc;
}

extern "C" int lb_composite_rgb_fwd(const float* weights, const float* colors, int n_rays, int n_samples, int n_colors,
                                    int white_bkgd, float* rgb, void* stream) {
    LB_REQUIRE(n_rays == 0 || (weights && colors && rgb), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1 && n_colors >= 1 && n_colors <= 4, "bad sizes (n_colors in [1, 4])");
    if (n_rays == 0) return LB_OK;
    composite_rgb_fwd_kernel<<<(n_rays + 7) / 8, 256, 0, (cudaStream_t)stream>>>(weights, colors, n_rays, n_samples,
                                                                               n_colors, white_bkgd, rgb);
    LB_LAUNCH_CHECK("lb_composite_rgb_fwd");
    return LB_OK;
}

extern "C" int lb_composite_rgb_bwd(const float* weights, const float* colors, const float* g_rgb, int n_rays,
                                    int n_samples, int n_colors, int white_bkgd, float* d_weights, float* d_colors,
                                    void* stream) {
    LB_REQUIRE(n_rays == 0 || (weights && colors && g_rgb), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(d_weights || d_colors, "nothing to compute");
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1 && n_colors >= 1 && n_colors <= 4, "bad sizes (n_colors in [1, 4])");
    if (n_rays == 0) return LB_OK;
    const size_t n = (size_t)n_rays * n_samples;
    composite_rgb_bwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        weights, colors, g_rgb, n_rays, n_samples, n_colors, white_bkgd, d_weights, d_colors);
    LB_LAUNCH_CHECK("lb_composite_rgb_bwd");
    return LB_OK;
}
