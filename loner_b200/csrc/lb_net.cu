// RGB branch, remaining pieces: spherical-harmonics direction encoding (forward / input gradient), the backward
// pass of the fused network (fp32 CUDA-core tile GEMMs; the tcgen05 forward is lb_net_umma.cu) and colour
// compositing.
//
// Replaces tcnn.Encoding(3, {"otype": "SphericalHarmonics", "degree": 4}) (reference src/models/nerf_tcnn.py:43,86),
// the autograd backward of tcnn.Network (nerf_tcnn.py:50-52,89-91) and the colour lines of raw2outputs
// (src/models/rendering_tcnn.py:284-289).
#include "lb_common.cuh"

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
// network backward: per 128-point tile, layer by layer from the output, two tile GEMMs per layer
//   dW_l[j,k] += sum_p dZ_l[p,j] Hprev[p,k]        dHprev[p,k] = sum_j dZ_l[p,j] W_l[j,k]
// dZ / Hprev / W_l tiles live in shared memory (row stride 68 floats); the weight-gradient accumulators of all
// layers stay in shared memory for the CTA's lifetime (every entry has one owner thread) and are flushed with
// one atomicAdd per entry at the end.
// ------------------------------------------------------------------------------------------
constexpr int kT = 128, kW = 64, kLd = 68, kOutPad = 16, kBwdThreads = 256;

struct NetShape {
    int n_in, n_hidden, n_out, out_act;
};

__device__ __forceinline__ size_t layer_param_off(const NetShape& s, int l) {
    return l == 0 ? 0 : (size_t)kW * s.n_in + (size_t)(l - 1) * kW * kW;
}

// activation tile written by lb_net_fwd: [16-byte chunk q of the 64-wide row][row r] float4 -> shared tile [r][q]
__device__ __forceinline__ void load_act_tile(const float* __restrict__ acts, size_t layer, size_t n_tiles, size_t tile,
                                              int valid_rows, float* dst) {
    const float4* src = reinterpret_cast<const float4*>(acts) + (layer * n_tiles + tile) * (size_t)(kT * 16);
    for (int it = threadIdx.x; it < kT * 16; it += kBwdThreads) {
        const int q = it / kT, r = it - q * kT;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < valid_rows) v = __ldg(src + it);
        *reinterpret_cast<float4*>(dst + r * kLd + q * 4) = v;
    }
}

// rows x cols fp32 (row-major, leading dimension ld_src) -> shared tile (leading dimension kLd); rows beyond
// `valid_rows` are zero
__device__ __forceinline__ void load_tile_f4(const float* __restrict__ src, size_t ld_src, int rows, int cols,
                                             int valid_rows, float* dst) {
    const int c4 = cols / 4;
    for (int it = threadIdx.x; it < rows * c4; it += kBwdThreads) {
        const int r = it / c4, c = it - r * c4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < valid_rows) v = __ldg(reinterpret_cast<const float4*>(src + (size_t)r * ld_src + c * 4));
        *reinterpret_cast<float4*>(dst + r * kLd + c * 4) = v;
    }
}

__global__ void __launch_bounds__(kBwdThreads, 1)
net_bwd_kernel(const float* __restrict__ params, const float* __restrict__ x, const float* __restrict__ acts,
               const float* __restrict__ out, const float* __restrict__ d_out, size_t n_points, NetShape s,
               float* __restrict__ grad_params, float* __restrict__ d_x) {
    extern __shared__ __align__(16) float smf[];
    float* sDZ = smf;                         // [128][68]
    float* sH = sDZ + kT * kLd;               // [128][68]
    float* sW = sH + kT * kLd;                // [64][68]
    float* sDO = sW + kW * kLd;               // [128][16]
    float* sAcc = sDO + kT * kOutPad;         // per layer [64][64] (+ output layer [16][64])
    const int tid = threadIdx.x;
    const int n_acc = s.n_hidden * kW * kW + kOutPad * kW;
    for (int i = tid; i < n_acc; i += kBwdThreads) sAcc[i] = 0.0f;
    const size_t out_off = layer_param_off(s, s.n_hidden);
    const int ty = tid >> 4, tx = tid & 15;   // 16 x 16 thread grid over a 64 x 64 (or 128 x 64) output tile

    const size_t n_tiles = (n_points + kT - 1) / kT;
    for (size_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const size_t p0 = tile * (size_t)kT;
        const int valid = (int)((n_points - p0) < (size_t)kT ? (n_points - p0) : (size_t)kT);
        __syncthreads();
        // output layer: dZo[p,c] = d_out (* o (1 - o) for the fused sigmoid)
        for (int it = tid; it < kT * kOutPad; it += kBwdThreads) {
            const int r = it / kOutPad, c = it - r * kOutPad;
            float g = 0.0f;
            if (r < valid && c < s.n_out) {
                g = __ldg(d_out + (p0 + r) * s.n_out + c);
                if (s.out_act == 1) {
                    const float o = __ldg(out + (p0 + r) * s.n_out + c);
                    g *= o * (1.0f - o);
                }
            }
            sDO[it] = g;
        }
        load_act_tile(acts, (size_t)(s.n_hidden - 1), n_tiles, tile, valid, sH);
        load_tile_f4(params + out_off, kW, kOutPad, kW, kOutPad, sW);
        __syncthreads();
        // dW_out[c, j] += sum_p dZo[p, c] H[p, j]
        for (int it = tid; it < s.n_out * kW; it += kBwdThreads) {
            const int c = it / kW, j = it - c * kW;
            float a = 0.0f;
#pragma unroll 8
            for (int p = 0; p < kT; ++p) a = fmaf(sDO[p * kOutPad + c], sH[p * kLd + j], a);
            sAcc[s.n_hidden * kW * kW + c * kW + j] += a;
        }
        // dZ_{last hidden}[p, j] = (sum_c dZo[p, c] Wout[c, j]) * [H[p, j] > 0]
        for (int it = tid; it < kT * kW; it += kBwdThreads) {
            const int p = it / kW, j = it - p * kW;
            float a = 0.0f;
            for (int c = 0; c < s.n_out; ++c) a = fmaf(sDO[p * kOutPad + c], sW[c * kLd + j], a);
            sDZ[p * kLd + j] = sH[p * kLd + j] > 0.0f ? a : 0.0f;
        }
        for (int l = s.n_hidden - 1; l >= 0; --l) {
            const int K = l == 0 ? s.n_in : kW;
            __syncthreads();                  // sDZ complete; sH / sW no longer read
            if (l > 0) load_act_tile(acts, (size_t)(l - 1), n_tiles, tile, valid, sH);
            else load_tile_f4(x + p0 * s.n_in, s.n_in, kT, K, valid, sH);
            load_tile_f4(params + layer_param_off(s, l), K, kW, K, kW, sW);
            __syncthreads();
            const int k0 = tx * 4;
            if (k0 < K) {
                // dW_l[j0..j0+3][k0..k0+3] += sum_p dZ[p][j] Hprev[p][k]
                const int j0 = ty * 4;
                float acc[4][4];
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 4; ++b) acc[a][b] = 0.0f;
#pragma unroll 4
                for (int p = 0; p < kT; ++p) {
                    const float4 dz = *reinterpret_cast<const float4*>(sDZ + p * kLd + j0);
                    const float4 h = *reinterpret_cast<const float4*>(sH + p * kLd + k0);
                    const float dzv[4] = {dz.x, dz.y, dz.z, dz.w}, hv[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
                    for (int a = 0; a < 4; ++a)
#pragma unroll
                        for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(dzv[a], hv[b], acc[a][b]);
                }
                float* dst = sAcc + l * kW * kW;
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 4; ++b) dst[(j0 + a) * kW + k0 + b] += acc[a][b];   // owner-exclusive entry
            }
            const bool need_prev = l > 0 || d_x != nullptr;
            float dprev[8][4];
            if (need_prev && k0 < K) {
                // dHprev[p0r..p0r+7][k0..k0+3] = sum_j dZ[p][j] W[j][k]
                const int r0 = ty * 8;
#pragma unroll
                for (int a = 0; a < 8; ++a)
#pragma unroll
                    for (int b = 0; b < 4; ++b) dprev[a][b] = 0.0f;
#pragma unroll 2
                for (int j4 = 0; j4 < kW; j4 += 4) {
                    float wv[4][4];
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        const float4 w = *reinterpret_cast<const float4*>(sW + (j4 + jj) * kLd + k0);
                        wv[jj][0] = w.x; wv[jj][1] = w.y; wv[jj][2] = w.z; wv[jj][3] = w.w;
                    }
#pragma unroll
                    for (int a = 0; a < 8; ++a) {
                        const float4 dz = *reinterpret_cast<const float4*>(sDZ + (r0 + a) * kLd + j4);
                        const float dzv[4] = {dz.x, dz.y, dz.z, dz.w};
#pragma unroll
                        for (int jj = 0; jj < 4; ++jj)
#pragma unroll
                            for (int b = 0; b < 4; ++b) dprev[a][b] = fmaf(dzv[jj], wv[jj][b], dprev[a][b]);
                    }
                }
            }
            __syncthreads();                  // every read of sDZ for this layer is done
            if (need_prev && k0 < K) {
                const int r0 = ty * 8;
#pragma unroll
                for (int a = 0; a < 8; ++a) {
                    if (l > 0) {
                        const float4 h = *reinterpret_cast<const float4*>(sH + (r0 + a) * kLd + k0);
                        *reinterpret_cast<float4*>(sDZ + (r0 + a) * kLd + k0) =
                            make_float4(h.x > 0.f ? dprev[a][0] : 0.f, h.y > 0.f ? dprev[a][1] : 0.f,
                                        h.z > 0.f ? dprev[a][2] : 0.f, h.w > 0.f ? dprev[a][3] : 0.f);
                    } else if (r0 + a < valid) {
                        *reinterpret_cast<float4*>(d_x + (p0 + r0 + a) * s.n_in + k0) =
                            make_float4(dprev[a][0], dprev[a][1], dprev[a][2], dprev[a][3]);
                    }
                }
            }
        }
    }
    __syncthreads();
    if (grad_params) {
        for (int l = 0; l < s.n_hidden; ++l) {
            const int K = l == 0 ? s.n_in : kW;
            const size_t off = layer_param_off(s, l);
            for (int i = tid; i < kW * K; i += kBwdThreads) {
                const int j = i / K, k = i - j * K;
                atomicAdd(grad_params + off + i, sAcc[l * kW * kW + j * kW + k]);
            }
        }
        for (int i = tid; i < s.n_out * kW; i += kBwdThreads)
            atomicAdd(grad_params + out_off + i, sAcc[s.n_hidden * kW * kW + i]);
    }
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
    LB_REQUIRE(d01 && out, "null pointer");
    LB_REQUIRE(n_points >= 0, "negative size");
    if (n_points == 0) return LB_OK;
    sh4_fwd_kernel<<<(unsigned)((n_points + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d01, (size_t)n_points, out);
    LB_LAUNCH_CHECK("lb_sh4_fwd");
    return LB_OK;
}

extern "C" int lb_sh4_bwd(const float* d01, const float* d_out, int64_t n_points, float* d_d01, void* stream) {
    LB_REQUIRE(d01 && d_out && d_d01, "null pointer");
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
    LB_REQUIRE(x && params && acts && d_out, "null pointer");
    LB_REQUIRE(grad_params || d_x, "nothing to compute");
    LB_REQUIRE(n_points >= 0, "negative size");
    LB_REQUIRE(n_in >= 16 && n_in <= 64 && n_in % 16 == 0, "n_in must be 16, 32, 48 or 64");
    LB_REQUIRE(n_hidden >= 1 && n_hidden <= 6, "n_hidden_layers must be in [1, 6]");
    LB_REQUIRE(n_out >= 1 && n_out <= kOutPad, "n_out must be in [1, 16]");
    LB_REQUIRE(out_act == 0 || (out_act == 1 && out), "out_act 1 (sigmoid) needs the forward output");
    if (n_points == 0) return LB_OK;
    const NetShape s{n_in, n_hidden, n_out, out_act};
    const size_t smem = sizeof(float) * ((size_t)2 * kT * kLd + kW * kLd + kT * kOutPad + (size_t)n_hidden * kW * kW + kOutPad * kW);
    if (smem > 227 * 1024) {
        lb_set_error("lb_net_bwd: %d hidden layers do not fit in shared memory", n_hidden);
        return LB_ERR_UNSUPPORTED;
    }
    int rc = lb_check_cuda(cudaFuncSetAttribute(net_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                           "net_bwd smem opt-in");
    if (rc) return rc;
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const size_t tiles = ((size_t)n_points + kT - 1) / kT;
    const int grid = (int)(tiles < (size_t)sms ? tiles : (size_t)sms);
    net_bwd_kernel<<<grid, kBwdThreads, smem, (cudaStream_t)stream>>>(params, x, acts, out, d_out, (size_t)n_points, s,
                                                                      grad_params, d_x);
    LB_LAUNCH_CHECK("lb_net_bwd");
    return LB_OK;
}

extern "C" int lb_composite_rgb_fwd(const float* weights, const float* colors, int n_rays, int n_samples, int n_colors,
                                    int white_bkgd, float* rgb, void* stream) {
    LB_REQUIRE(weights && colors && rgb, "null pointer");
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
    LB_REQUIRE(weights && colors && g_rgb, "null pointer");
    LB_REQUIRE(d_weights || d_colors, "nothing to compute");
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1 && n_colors >= 1 && n_colors <= 4, "bad sizes (n_colors in [1, 4])");
    if (n_rays == 0) return LB_OK;
    const size_t n = (size_t)n_rays * n_samples;
    composite_rgb_bwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
        weights, colors, g_rgb, n_rays, n_samples, n_colors, white_bkgd, d_weights, d_colors);
    LB_LAUNCH_CHECK("lb_composite_rgb_bwd");
    return LB_OK;
}
