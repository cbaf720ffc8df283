// R5/R6 + L1-L5: alpha compositing, JS dynamic-margin depth loss and their analytic backward.
//
// Replaces raw2outputs / raw2outputs_adjusted (reference src/models/rendering_tcnn.py:219-295, 70-214),
// the lidar branch of Optimizer.compute_loss (src/mapping/optimizer.py:719-848), get_weights_gt
// (src/models/losses.py:29-51) and the autograd replay of all of them.
//
// One warp per ray.  A ray is walked in chunks of 128 samples: each lane owns 4 consecutive samples
// (one coalesced float4 per array), multiplies its transmittance factors serially, and a 5-step
// shuffle scan stitches the lanes; a carry stitches the chunks.  The backward walks the chunks in
// reverse with a suffix scan.  Per-sample data are re-read from L1/L2 between passes instead of
// being held in registers, so any S up to 8192 works with the same code.
#include "lb_common.cuh"

namespace {

constexpr int kWarps = 4;            // rays per block
constexpr int kE = 4;                // samples per lane per chunk
constexpr int kChunk = 32 * kE;      // samples per chunk
constexpr int kMaxChunks = 64;       // S <= 8192

struct Ray {
    float dnorm, far, dx, dy, dz;
};

__device__ __forceinline__ Ray load_ray(const float* __restrict__ rays, int r) {
    const float* p = rays + (size_t)r * LB_RAY_STRIDE;
    Ray h;
    h.dx = __ldg(p + 3); h.dy = __ldg(p + 4); h.dz = __ldg(p + 5);
    h.dnorm = sqrtf(h.dx * h.dx + h.dy * h.dy + h.dz * h.dz);
    h.far = __ldg(p + 12);
    return h;
}

__device__ __forceinline__ void load4(const float* __restrict__ row, int i0, int S, bool vec, float fill, float* v) {
    if (vec && i0 + 3 < S) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(row + i0));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
        for (int e = 0; e < kE; ++e) v[e] = (i0 + e < S) ? __ldg(row + i0 + e) : fill;
    }
}

__device__ __forceinline__ void store4(float* __restrict__ row, int i0, int S, bool vec, const float* v) {
    if (vec && i0 + 3 < S) {
        *reinterpret_cast<float4*>(row + i0) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
        for (int e = 0; e < kE; ++e)
            if (i0 + e < S) row[i0 + e] = v[e];
    }
}

// Per-chunk sample quantities for this lane's 4 samples.
struct Chunk {
    float z[kE], dz[kE], r[kE], ex[kE], alpha[kE], q[kE], T[kE], w[kE];
    bool live[kE], pos[kE];
};

// Computes alpha, q, transmittance T (exclusive product, incl. carry-in) and weights for one chunk.
// Returns the carry-out (product of all q up to the end of the chunk).
__device__ __forceinline__ float chunk_forward(const float* __restrict__ sig, const float* __restrict__ zr,
                                               const float* __restrict__ nz, int S, bool vec, int c, int lane,
                                               const Ray& ray, float carry, Chunk& k) {
    const int i0 = c * kChunk + lane * kE;
    float s[kE], n[kE];
    load4(sig, i0, S, vec, 0.0f, s);
    load4(zr, i0, S, vec, 0.0f, k.z);
    if (nz) load4(nz, i0, S, vec, 0.0f, n);
    const float znext = (i0 + kE < S) ? __ldg(zr + i0 + kE) : 0.0f;
    float prod = 1.0f;
#pragma unroll
    for (int e = 0; e < kE; ++e) {
        const int i = i0 + e;
        k.live[e] = i < S;
        const float zn = (e + 1 < kE) ? k.z[e + 1] : znext;
        k.dz[e] = (i < S - 1) ? (zn - k.z[e]) : 1e10f;
        const float delta = k.dz[e] * ray.dnorm;
        const float sv = nz ? s[e] + n[e] : s[e];
        k.pos[e] = sv > 0.0f;
        k.r[e] = fmaxf(sv, 0.0f);
        k.ex[e] = expf(-delta * k.r[e]);
        k.alpha[e] = 1.0f - k.ex[e];
        k.q[e] = k.live[e] ? (1.0f - k.alpha[e] + 1e-10f) : 1.0f;
        k.T[e] = prod;          // exclusive within the lane
        prod *= k.q[e];
    }
    // exclusive scan of lane products across the warp
    float incl = prod;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const float t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl *= t;
    }
    float excl = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) excl = 1.0f;
    const float base = carry * excl;
#pragma unroll
    for (int e = 0; e < kE; ++e) {
        k.T[e] *= base;
        k.w[e] = k.live[e] ? k.alpha[e] * k.T[e] : 0.0f;
    }
    return carry * __shfl_sync(0xffffffffu, incl, 31);
}

struct RaySums {
    float A, wz;   // sum w, sum w*z
};

// Forward compositing pass: writes weights, returns A and sum(w z); records chunk carries if asked.
__device__ __forceinline__ RaySums composite_pass(const float* __restrict__ sig, const float* __restrict__ zr,
                                                  const float* __restrict__ nz, int S, bool vec, int lane,
                                                  const Ray& ray, float* __restrict__ w_out, float* carries) {
    const int n_chunks = (S + kChunk - 1) / kChunk;
    float carry = 1.0f, A = 0.0f, wz = 0.0f;
    for (int c = 0; c < n_chunks; ++c) {
        if (carries && lane == 0) carries[c] = carry;
        Chunk k;
        carry = chunk_forward(sig, zr, nz, S, vec, c, lane, ray, carry, k);
#pragma unroll
        for (int e = 0; e < kE; ++e) { A += k.w[e]; wz = fmaf(k.w[e], k.z[e], wz); }
        if (w_out) store4(w_out, c * kChunk + lane * kE, S, vec, k.w);
    }
    RaySums s;
    s.A = warp_sum(A);
    s.wz = warp_sum(wz);
    return s;
}

// sum_i w_i (c - z_i)^2 re-reading w and z
__device__ __forceinline__ float second_moment(const float* __restrict__ w_row, const float* __restrict__ zr,
                                               int S, bool vec, int lane, float centre, float zscale) {
    float acc = 0.0f;
    for (int i0 = lane * kE; i0 < S; i0 += kChunk) {
        float w[kE], z[kE];
        load4(w_row, i0, S, vec, 0.0f, w);
        load4(zr, i0, S, vec, 0.0f, z);
#pragma unroll
        for (int e = 0; e < kE; ++e) { const float d = centre - z[e] * zscale; acc = fmaf(w[e], d * d, acc); }
    }
    return warp_sum(acc);
}

// both second moments the loss needs in one sweep over (w, z): about D in cube units (the variance output) and about
// the weighted mean in metres (the JS score).  Same operations per accumulator as two second_moment calls.
__device__ __forceinline__ float2 second_moment2(const float* __restrict__ w_row, const float* __restrict__ zr, int S,
                                                 bool vec, int lane, float c0, float c1, float zscale1) {
    float a0 = 0.0f, a1 = 0.0f;
    for (int i0 = lane * kE; i0 < S; i0 += kChunk) {
        float w[kE], z[kE];
        load4(w_row, i0, S, vec, 0.0f, w);
        load4(zr, i0, S, vec, 0.0f, z);
#pragma unroll
        for (int e = 0; e < kE; ++e) {
            const float d0 = c0 - z[e] * 1.0f, d1 = c1 - z[e] * zscale1;
            a0 = fmaf(w[e], d0 * d0, a0);
            a1 = fmaf(w[e], d1 * d1, a1);
        }
    }
    return make_float2(warp_sum(a0), warp_sum(a1));
}

// Reverse pass: given g_i = dL/dw_i (functor), produce d_sigma and the |d| / far partials.
//   B_k = sum_{i>k} g_i w_i ;  dL/dalpha_k = g_k T_k - B_k / q_k
//   dL/dsigma_k = dalpha_k * delta_k * exp(-delta_k r_k) * [sigma_k + n_k > 0]
//   dL/d|d|    += dalpha_k * r_k * exp(-delta_k r_k) * dz_k
template <typename G>
__device__ __forceinline__ float reverse_pass(const float* __restrict__ sig, const float* __restrict__ zr,
                                              const float* __restrict__ nz, int S, bool vec, int lane,
                                              const Ray& ray, const float* carries, float* __restrict__ ds_out, G g_of) {
    const int n_chunks = (S + kChunk - 1) / kChunk;
    float Bcarry = 0.0f, dnorm_acc = 0.0f;
    for (int c = n_chunks - 1; c >= 0; --c) {
        Chunk k;
        chunk_forward(sig, zr, nz, S, vec, c, lane, ray, carries[c], k);
        const int i0 = c * kChunk + lane * kE;
        float g[kE], gw[kE];
        float lane_sum = 0.0f;
#pragma unroll
        for (int e = 0; e < kE; ++e) {
            g[e] = k.live[e] ? g_of(i0 + e, k.w[e], k.z[e]) : 0.0f;
            gw[e] = g[e] * k.w[e];
            lane_sum += gw[e];
        }
        // suffix (exclusive) sum across lanes: sum over lanes > lane
        float incl = lane_sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float t = __shfl_down_sync(0xffffffffu, incl, o);
            if (lane + o < 32) incl += t;
        }
        float after = __shfl_down_sync(0xffffffffu, incl, 1);
        if (lane == 31) after = 0.0f;
        float B = Bcarry + after;           // contributions of all samples after this lane's segment
        float ds[kE];
#pragma unroll
        for (int e = kE - 1; e >= 0; --e) {
            const float dalpha = g[e] * k.T[e] - B / k.q[e];
            const float delta = k.dz[e] * ray.dnorm;
            ds[e] = (k.live[e] && k.pos[e]) ? dalpha * delta * k.ex[e] : 0.0f;
            if (k.live[e]) dnorm_acc = fmaf(dalpha * k.r[e] * k.ex[e], k.dz[e], dnorm_acc);
            B += gw[e];
        }
        if (ds_out) store4(ds_out, i0, S, vec, ds);
        Bcarry += __shfl_sync(0xffffffffu, incl, 0);
    }
    return warp_sum(dnorm_acc);
}

__device__ __forceinline__ float sgn(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }

// KL(N(m1,s1) || N(m2,s2))  (optimizer.py:913-920)
__device__ __forceinline__ float kl_gauss(float m1, float s1, float m2, float s2) {
    const float d = m1 - m2;
    return logf(s2 / s1) + (s1 * s1 + d * d) / (2.0f * s2 * s2) - 0.5f;
}

// ------------------------------------------------------------------------------------------
// raw2outputs forward ('default' and 'adjusted')
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarps * 32)
composite_fwd_kernel(const float* __restrict__ sigma, const float* __restrict__ zv, const float* __restrict__ rays,
                     const float* __restrict__ noise, int n_rays, int S, int strategy, float* __restrict__ weights,
                     float* __restrict__ depth, float* __restrict__ opacity, float* __restrict__ variance) {
    const int lane = threadIdx.x & 31;
    const int r = blockIdx.x * kWarps + (threadIdx.x >> 5);
    if (r >= n_rays) return;
    const bool vec = (S % 4) == 0;
    const Ray ray = load_ray(rays, r);
    const float* sig = sigma + (size_t)r * S;
    const float* zr = zv + (size_t)r * S;
    const float* nz = (noise && strategy == 0) ? noise + (size_t)r * S : nullptr;
    float* w_row = weights + (size_t)r * S;
    float D;
    RaySums s;
    if (strategy == 0) {
        s = composite_pass(sig, zr, nz, S, vec, lane, ray, w_row, nullptr);
        D = s.wz + (1.0f - s.A) * ray.far;           // rendering_tcnn.py:274-278
    } else {
        // median-transmittance depth: z_i at the unique i >= 1 with T_i <= 0.5 < T_{i-1}
        // (rendering_tcnn.py:121-132); 0 if the transmittance never crosses.
        const int n_chunks = (S + kChunk - 1) / kChunk;
        float carry = 1.0f, A = 0.0f, wz = 0.0f, hit = 0.0f, t_last = 1.0f;
        for (int c = 0; c < n_chunks; ++c) {
            Chunk k;
            carry = chunk_forward(sig, zr, nullptr, S, vec, c, lane, ray, carry, k);
            // T of the sample just before this lane's first one
            float t_prev = __shfl_up_sync(0xffffffffu, k.T[kE - 1], 1);
            if (lane == 0) t_prev = t_last;
#pragma unroll
            for (int e = 0; e < kE; ++e) {
                const int i = c * kChunk + lane * kE + e;
                if (k.live[e] && i >= 1 && !(k.T[e] > 0.5f) && (t_prev > 0.5f)) hit += k.z[e];
                t_prev = k.T[e];
                A += k.w[e]; wz = fmaf(k.w[e], k.z[e], wz);
            }
            t_last = __shfl_sync(0xffffffffu, k.T[kE - 1], 31);
            store4(w_row, c * kChunk + lane * kE, S, vec, k.w);
        }
        s.A = warp_sum(A); s.wz = warp_sum(wz);
        D = warp_sum(hit);
    }
    float V = 0.0f;
    if (variance) {
        __syncwarp();
        V = second_moment(w_row, zr, S, vec, lane, D, 1.0f);   // rendering_tcnn.py:292 / :211
    }
    if (lane == 0) {
        depth[r] = D;
        opacity[r] = s.A;
        if (variance) variance[r] = V;
    }
}

// ------------------------------------------------------------------------------------------
// raw2outputs backward for arbitrary upstream gradients
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarps * 32)
composite_bwd_kernel(const float* __restrict__ sigma, const float* __restrict__ zv, const float* __restrict__ rays,
                     const float* __restrict__ noise, const float* __restrict__ g_w, const float* __restrict__ g_depth,
                     const float* __restrict__ g_opacity, const float* __restrict__ g_variance, int n_rays, int S,
                     float* __restrict__ d_sigma, float* __restrict__ d_rays) {
    __shared__ float carries_s[kWarps][kMaxChunks];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r = blockIdx.x * kWarps + warp;
    if (r >= n_rays) return;
    const bool vec = (S % 4) == 0;
    const Ray ray = load_ray(rays, r);
    const float* sig = sigma + (size_t)r * S;
    const float* zr = zv + (size_t)r * S;
    const float* nz = noise ? noise + (size_t)r * S : nullptr;
    float* carries = carries_s[warp];
    const RaySums s = composite_pass(sig, zr, nz, S, vec, lane, ray, nullptr, carries);
    __syncwarp();
    const float D = s.wz + (1.0f - s.A) * ray.far;
    const float gd = g_depth ? __ldg(g_depth + r) : 0.0f;
    const float go = g_opacity ? __ldg(g_opacity + r) : 0.0f;
    const float gv = g_variance ? __ldg(g_variance + r) : 0.0f;
    const float dV_dD = 2.0f * (D * s.A - s.wz);          // d/dD sum w (D - z)^2
    const float gD = gd + gv * dV_dD;                     // total gradient reaching D
    const float* gw_row = g_w ? g_w + (size_t)r * S : nullptr;
    const float far = ray.far;
    auto g_of = [&](int i, float w, float z) {
        const float dz = D - z;
        return (gw_row ? __ldg(gw_row + i) : 0.0f) + gD * (z - far) + go + gv * dz * dz;
    };
    const float dnorm = reverse_pass(sig, zr, nz, S, vec, lane, ray, carries, d_sigma + (size_t)r * S, g_of);
    if (d_rays && lane == 0) {
        float* dr = d_rays + (size_t)r * LB_RAY_STRIDE;
        const float inv = ray.dnorm > 0.0f ? 1.0f / ray.dnorm : 0.0f;
        atomicAdd(dr + 3, dnorm * ray.dx * inv);
        atomicAdd(dr + 4, dnorm * ray.dy * inv);
        atomicAdd(dr + 5, dnorm * ray.dz * inv);
        atomicAdd(dr + 12, gD * (1.0f - s.A));
    }
}

// ------------------------------------------------------------------------------------------
// fused compositing + loss + analytic gradient
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kWarps * 32)
composite_loss_kernel(const float* __restrict__ sigma, const float* __restrict__ zv, const float* __restrict__ rays,
                      const float* __restrict__ noise, const float* __restrict__ depth_gt, lb_loss_config cfg,
                      const float* __restrict__ norm_dev, int n_rays, int S, float* __restrict__ weights, float* __restrict__ depth,
                      float* __restrict__ opacity, float* __restrict__ variance, float* __restrict__ js_out,
                      float* __restrict__ eps_out, float* __restrict__ loss_parts, float* __restrict__ d_sigma,
                      float* __restrict__ d_rays) {
    __shared__ float carries_s[kWarps][kMaxChunks];
    __shared__ float parts_s[kWarps][4];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int r = blockIdx.x * kWarps + warp;
    const bool active = r < n_rays;
    float part[4] = {0.f, 0.f, 0.f, 0.f};
    if (active) {
        const bool vec = (S % 4) == 0;
        const Ray ray = load_ray(rays, r);
        const float* sig = sigma + (size_t)r * S;
        const float* zr = zv + (size_t)r * S;
        const float* nz = noise ? noise + (size_t)r * S : nullptr;
        float* w_row = weights + (size_t)r * S;
        float* carries = carries_s[warp];
        const float sc = cfg.scale_factor;

        // pass 1: weights, opacity, expected depth
        const RaySums s = composite_pass(sig, zr, nz, S, vec, lane, ray, w_row, carries);
        __syncwarp();
        const float A = s.A;
        const float D = s.wz + (1.0f - A) * ray.far;
        // depth_gt = NaN marks a MASKED ray (lb_build_lidar_rays, mask_span): it is rendered like any other ray (the outputs
        // below are defined) but contributes no loss term and no gradient, and lb_count_opaque leaves it out of both
        // normalisers -- the arithmetic of a batch from which the reference had dropped the ray
        const float dg_in = __ldg(depth_gt + r);
        const bool masked = dg_in != dg_in;
        const float dg = masked ? 0.0f : dg_in;
        const float g = dg * sc;
        const bool opaque = (dg > 0.0f) && !(dg > ray.far);        // optimizer.py:724-727

        // pass 2: variance about D (output) and std about the weighted mean (JS score)
        const float mean = (s.wz * sc) / (A + 1e-10f);             // optimizer.py:739
        const float2 m2 = second_moment2(w_row, zr, S, vec, lane, D, mean, sc);
        const float V = m2.x;
        const float var = m2.y / (A + 1e-10f) + 1e-10f;
        const float sd = sqrtf(var);

        // JS divergence between N(g, (eps_min/3)^2) and N(mean, sd^2)  (optimizer.py:744-746, 922-925)
        const float eps_min = cfg.min_depth_eps;
        const float s1 = eps_min / 3.0f;
        const float mm = 0.5f * (g + mean);
        const float sm = 0.5f * sqrtf(s1 * s1 + sd * sd);
        float js = 0.5f * kl_gauss(g, s1, mm, sm) + 0.5f * kl_gauss(mean, sd, mm, sm);
        float eps;
        if (cfg.kind == LB_LOSS_L1_JS || cfg.kind == LB_LOSS_L2_JS) {
            if (js < cfg.min_js_score) js = 0.0f;                  // optimizer.py:763-765
            if (js > cfg.max_js_score) js = cfg.max_js_score;
            eps = eps_min * (1.0f + cfg.js_alpha * js);
        } else {
            eps = cfg.depth_eps;
        }
        const bool l2 = (cfg.kind == LB_LOSS_L2_JS || cfg.kind == LB_LOSS_L2_LOS);

        // pass 3: normaliser of the truncated-Gaussian target weights (losses.py:29-51)
        const float st = eps / 9.0f;
        const float inv_st = 1.0f / st;
        const float lo = g - eps, hi = g + eps;
        const float clipn = 0.5f * (1.0f + erff((eps / st) * 0.70710678118654752f)) -
                            0.5f * (1.0f + erff((-eps / st) * 0.70710678118654752f));
        auto w_target = [&](float z) {
            const float zm = z * sc;
            const float t = (zm - g) * inv_st;
            const float pdf = 0.3989422804014327f * expf(-0.5f * t * t);
            return (zm > lo && zm < hi) ? pdf * inv_st / clipn : 0.0f;
        };
        float wg_sum = 0.0f;
        if (opaque) {
            float acc = 0.0f;
            for (int i0 = lane * kE; i0 < S; i0 += kChunk) {
                float z[kE];
                load4(zr, i0, S, vec, 0.0f, z);
#pragma unroll
                for (int e = 0; e < kE; ++e)
                    if (i0 + e < S) acc += w_target(z[e]);
            }
            wg_sum = warp_sum(acc);
        }
        const float inv_wg = 1.0f / (wg_sum + 1e-6f);

        // pass 4 (reverse): line-of-sight term, dL/dw and the chain to sigma, |d|, far
        // global normalisers: {opaque-ray count, ray count} over all chunks and ranks, read from the device so that
        // neither costs a host round trip (lb_count_opaque + one all-reduce)
        const float n_op = norm_dev ? __ldg(norm_dev) : cfg.n_opaque;
        const float n_all = norm_dev ? __ldg(norm_dev + 1) : 0.0f;
        const float n_total = n_all > 0.0f ? n_all * (float)S : cfg.n_total_samples;
        const float inv_nop = (opaque && n_op > 0.0f) ? 1.0f / n_op : 0.0f;
        const float c_los = masked ? 0.0f : cfg.los_lambda / n_total;
        const float derr = sc * D - g;
        const float gD = inv_nop * 2.0f * cfg.depthloss_lambda * sc * derr;     // dL/dD
        const float gA = inv_nop * sgn(A - 1.0f);                                 // dL/dA
        const float far = ray.far;
        float los_acc = 0.0f;
        auto g_of = [&](int i, float w, float z) {
            const float wt = opaque ? w_target(z) * inv_wg : 0.0f;
            const float diff = w - wt;
            los_acc += l2 ? diff * diff : fabsf(diff);
            return c_los * (l2 ? 2.0f * diff : sgn(diff)) + gD * (z - far) + gA;
        };
        const float dnorm = reverse_pass(sig, zr, nz, S, vec, lane, ray, carries, d_sigma ? d_sigma + (size_t)r * S : nullptr, g_of);
        const float los = warp_sum(los_acc);
        if (lane == 0) {
            depth[r] = D;
            opacity[r] = A;
            if (variance) variance[r] = V;
            if (js_out) js_out[r] = js;
            if (eps_out) eps_out[r] = eps;
            if (d_rays) {
                float* dr = d_rays + (size_t)r * LB_RAY_STRIDE;
                const float inv = ray.dnorm > 0.0f ? 1.0f / ray.dnorm : 0.0f;
#pragma unroll
                for (int i = 0; i < LB_RAY_STRIDE; ++i) dr[i] = 0.0f;
                dr[3] = dnorm * ray.dx * inv;
                dr[4] = dnorm * ray.dy * inv;
                dr[5] = dnorm * ray.dz * inv;
                dr[12] = gD * (1.0f - A);
            }
            part[0] = opaque ? derr * derr : 0.0f;
            part[1] = masked ? 0.0f : los;
            part[2] = opaque ? fabsf(A - 1.0f) : 0.0f;
            part[3] = opaque ? 1.0f : 0.0f;
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) parts_s[warp][i] = part[i];
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        float t = 0.0f;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) t += parts_s[w][threadIdx.x];
        atomicAdd(loss_parts + threadIdx.x, t);
    }
}

__global__ void count_opaque_kernel(const float* __restrict__ rays, const float* __restrict__ depth_gt, int n_rays,
                                    float* __restrict__ count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool op = false, alive = false;
    if (i < n_rays) {
        const float dg = __ldg(depth_gt + i);
        alive = dg == dg;
        op = (dg > 0.0f) && !(dg > __ldg(rays + (size_t)i * LB_RAY_STRIDE + 12));
    }
    const unsigned m = __ballot_sync(0xffffffffu, op);
    const unsigned a = __ballot_sync(0xffffffffu, alive);
    if ((threadIdx.x & 31) == 0) {
        if (m) atomicAdd(count, (float)__popc(m));
        if (a) atomicAdd(count + 1, (float)__popc(a));       // rays of the batch that are not masked (depth_gt = NaN)
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------
// C-ABI
// ------------------------------------------------------------------------------------------
static int check_rs(int n_rays, int n_samples) {
    if (n_rays < 0 || n_samples < 2 || n_samples > kChunk * kMaxChunks) {
        lb_set_error("composite: need n_rays >= 0 and 2 <= n_samples <= %d", kChunk * kMaxChunks);
        return LB_ERR_BAD_ARGUMENT;
    }
    return LB_OK;
}

extern "C" int lb_composite_fwd(const float* sigma, const float* z, const float* rays, const float* noise,
                                int n_rays, int n_samples, int strategy, float* weights, float* depth,
                                float* opacity, float* variance, void* stream) {
    LB_REQUIRE(n_rays == 0 || (sigma && z && rays && weights && depth && opacity), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(strategy == 0 || strategy == 1, "strategy must be 0 (default) or 1 (adjusted)");
    if (int rc = check_rs(n_rays, n_samples)) return rc;
    if (n_rays == 0) return LB_OK;
    composite_fwd_kernel<<<(n_rays + kWarps - 1) / kWarps, kWarps * 32, 0, (cudaStream_t)stream>>>(
        sigma, z, rays, noise, n_rays, n_samples, strategy, weights, depth, opacity, variance);
    LB_LAUNCH_CHECK("lb_composite_fwd");
    return LB_OK;
}

extern "C" int lb_composite_bwd(const float* sigma, const float* z, const float* rays, const float* noise,
                                const float* g_w, const float* g_depth, const float* g_opacity,
                                const float* g_variance, int n_rays, int n_samples, float* d_sigma,
                                float* d_rays, void* stream) {
    LB_REQUIRE(n_rays == 0 || (sigma && z && rays && d_sigma), "null pointer");   // an empty batch may pass NULL buffers
    if (int rc = check_rs(n_rays, n_samples)) return rc;
    if (n_rays == 0) return LB_OK;
    composite_bwd_kernel<<<(n_rays + kWarps - 1) / kWarps, kWarps * 32, 0, (cudaStream_t)stream>>>(
        sigma, z, rays, noise, g_w, g_depth, g_opacity, g_variance, n_rays, n_samples, d_sigma, d_rays);
    LB_LAUNCH_CHECK("lb_composite_bwd");
    return LB_OK;
}

extern "C" int lb_composite_loss(const float* sigma, const float* z, const float* rays, const float* noise,
                                 const float* depth_gt, const lb_loss_config* cfg, const float* norm_dev,
                                 int n_rays, int n_samples,
                                 float* weights, float* depth, float* opacity, float* variance, float* js,
                                 float* eps, float* loss_parts, float* d_sigma, float* d_rays, void* stream) {
    LB_REQUIRE(cfg, "null pointer");
    LB_REQUIRE(n_rays == 0 || (sigma && z && rays && depth_gt && weights && depth && opacity && loss_parts), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(cfg->kind >= LB_LOSS_L1_JS && cfg->kind <= LB_LOSS_L2_LOS, "unknown loss kind");
    LB_REQUIRE(cfg->scale_factor > 0.0f && cfg->min_depth_eps > 0.0f, "scale_factor and min_depth_eps must be positive");
    if (int rc = check_rs(n_rays, n_samples)) return rc;
    if (n_rays == 0) return LB_OK;
    lb_loss_config c = *cfg;
    if (!(c.n_total_samples > 0.0f)) c.n_total_samples = (float)n_rays * (float)n_samples;
    LB_REQUIRE(norm_dev || c.n_opaque >= 0.0f, "n_opaque must be supplied (lb_count_opaque) and non-negative");
    composite_loss_kernel<<<(n_rays + kWarps - 1) / kWarps, kWarps * 32, 0, (cudaStream_t)stream>>>(
        sigma, z, rays, noise, depth_gt, c, norm_dev, n_rays, n_samples, weights, depth, opacity, variance, js, eps,
        loss_parts, d_sigma, d_rays);
    LB_LAUNCH_CHECK("lb_composite_loss");
    return LB_OK;
}

extern "C" int lb_count_opaque(const float* rays, const float* depth_gt, int n_rays, float* count, void* stream) {
    LB_REQUIRE(n_rays == 0 || (rays && depth_gt && count), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_rays >= 0, "negative size");
    if (n_rays == 0) return LB_OK;
    count_opaque_kernel<<<(n_rays + 255) / 256, 256, 0, (cudaStream_t)stream>>>(rays, depth_gt, n_rays, count);
    LB_LAUNCH_CHECK("lb_count_opaque");
    return LB_OK;
}
