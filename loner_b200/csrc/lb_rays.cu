// R0: LiDAR ray construction on the device, forward and backward.
//
// Replaces LidarRayDirections.build_lidar_rays (reference src/common/ray_utils.py:269-322) with get_far_val
// (ray_utils.py:31-60) as called by KeyFrame.build_lidar_rays (src/mapping/keyframe.py:75-105), fused with the
// index selection directions[:, idx] / distances[idx] (src/mapping/optimizer.py:389-400).  The pose enters as
// the rotation matrix R [3,3] and the ray origin o = (t + shift) / scale [3], both tiny device tensors that the
// host computes from the pose 6-vector with autograd (pose_utils.py:354-368), so the chain rule to the pose
// stays in PyTorch while the per-ray work (and its reduction to dL/dR, dL/do) runs here.
//
//   u = R dir;  d = u / |u|;  near = r0 / s;  far = min(r1 / s, exit(o, d))
//   exit(o, d) = min_axis max(clamp((-1 - o_a) / (d_a + 1e-15), 0), clamp((1 - o_a) / (d_a + 1e-15), 0))
//   rays[i] = [o, d, -d, 0, 0, near, far];  depths[i] = distance / s
#include "lb_common.cuh"

namespace {

constexpr int kThreads = 256;

struct Pose {
    float r[9], o[3];
};

__device__ __forceinline__ Pose load_pose(const float* __restrict__ rot, const float* __restrict__ origin) {
    Pose p;
#pragma unroll
    for (int i = 0; i < 9; ++i) p.r[i] = __ldg(rot + i);
#pragma unroll
    for (int i = 0; i < 3; ++i) p.o[i] = __ldg(origin + i);
    return p;
}

struct RayGeom {
    float u[3], inv_norm, d[3];
    float exit_t;      // cube exit distance
    int exit_axis;     // axis whose slab gives the exit, -1 when the clamp at 0 is active
    float exit_den;    // d_a + 1e-15 of that axis
};

__device__ __forceinline__ RayGeom ray_geometry(const Pose& p, float dx, float dy, float dz) {
    RayGeom g;
#pragma unroll
    for (int a = 0; a < 3; ++a) g.u[a] = p.r[3 * a] * dx + p.r[3 * a + 1] * dy + p.r[3 * a + 2] * dz;
    const float n = sqrtf(g.u[0] * g.u[0] + g.u[1] * g.u[1] + g.u[2] * g.u[2]);
    g.inv_norm = 1.0f / n;
#pragma unroll
    for (int a = 0; a < 3; ++a) g.d[a] = g.u[a] / n;
    g.exit_t = 3.402823466e38f;
    g.exit_axis = -1;
    g.exit_den = 1.0f;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float den = g.d[a] + 1e-15f;
        const float t0 = (-1.0f - p.o[a]) / den, t1 = (1.0f - p.o[a]) / den;
        const float m = fmaxf(fmaxf(t0, 0.0f), fmaxf(t1, 0.0f));
        if (m < g.exit_t) {
            g.exit_t = m;
            g.exit_axis = (m > 0.0f) ? a : -1;
            g.exit_den = den;
        }
    }
    return g;
}

__global__ void __launch_bounds__(kThreads)
build_rays_kernel(const float* __restrict__ directions, const float* __restrict__ distances,
                  const long long* __restrict__ indices, int n, int m, const float* __restrict__ rot,
                  const float* __restrict__ origin, float near, float far_max, float inv_scale, float const_depth,
                  float mask_span, float* __restrict__ rays, float* __restrict__ depths) {
    const int i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= m) return;
    const Pose p = load_pose(rot, origin);
    long long j = indices ? indices[i] : (long long)i;
    if (j < 0) j += n;                                   // torch-style negative index
    j = j < 0 ? 0 : (j >= n ? (long long)n - 1 : j);       // never read out of bounds
    const float dx = __ldg(directions + j), dy = __ldg(directions + (size_t)n + j), dz = __ldg(directions + 2 * (size_t)n + j);
    const RayGeom g = ray_geometry(p, dx, dy, dz);
    float* r = rays + (size_t)i * LB_RAY_STRIDE;
    r[0] = p.o[0]; r[1] = p.o[1]; r[2] = p.o[2];
    r[3] = g.d[0]; r[4] = g.d[1]; r[5] = g.d[2];
    r[6] = -g.d[0]; r[7] = -g.d[1]; r[8] = -g.d[2];
    r[9] = 0.0f; r[10] = 0.0f;
    r[11] = near;
    const float far = fminf(far_max, g.exit_t);
    r[12] = far;
    // The reference DROPS rays that leave the world cube within one metre of `near` (ray_utils.py:321-322: a
    // boolean-mask gather, i.e. a host sync and a data-dependent batch shape).  With mask_span > 0 such a ray stays in
    // the batch and is MASKED instead: depth = NaN, which lb_count_opaque / lb_composite_loss / lb_occ_grad read as
    // "this ray does not exist" (no loss term, no gradient, not counted in the normalisers).
    const float depth = distances ? __ldg(distances + j) * inv_scale : const_depth;
    depths[i] = (mask_span > 0.0f && !(far > near + mask_span)) ? __int_as_float(0x7fc00000) : depth;
}

// d_rot [9] += sum_i g_u (x) dir_i,  d_origin [3] += sum_i (g_o + far term)
__global__ void __launch_bounds__(kThreads)
build_rays_bwd_kernel(const float* __restrict__ directions, const long long* __restrict__ indices, int n, int m,
                      const float* __restrict__ rot, const float* __restrict__ origin, float far_max,
                      const float* __restrict__ d_rays, float* __restrict__ d_rot, float* __restrict__ d_origin) {
    __shared__ float red[kThreads / 32][12];
    const Pose p = load_pose(rot, origin);
    float acc[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) acc[k] = 0.0f;
    for (int i = blockIdx.x * kThreads + threadIdx.x; i < m; i += gridDim.x * kThreads) {
        long long j = indices ? indices[i] : (long long)i;
        j = j < 0 ? 0 : (j >= n ? (long long)n - 1 : j);
    j = j < 0 ? 0 : (j >= n ? (long long)n - 1 : j);       // never read out of bounds (the host wrapper validates)
        if (j < 0) j += n;
        const float dir[3] = {__ldg(directions + j), __ldg(directions + (size_t)n + j), __ldg(directions + 2 * (size_t)n + j)};
        const RayGeom g = ray_geometry(p, dir[0], dir[1], dir[2]);
        const float* gr = d_rays + (size_t)i * LB_RAY_STRIDE;
        float go[3], gd[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            go[a] = __ldg(gr + a);
            gd[a] = __ldg(gr + 3 + a) - __ldg(gr + 6 + a);       // view direction = -d
        }
        const float gfar = __ldg(gr + 12);
        if (g.exit_t < far_max && g.exit_axis >= 0) {            // far = exit(o, d): (+-1 - o_a) / (d_a + eps)
#pragma unroll
            for (int a = 0; a < 3; ++a) {
                if (a == g.exit_axis) {
                    go[a] -= gfar / g.exit_den;
                    gd[a] -= gfar * g.exit_t / g.exit_den;
                }
            }
        }
        const float dot = g.d[0] * gd[0] + g.d[1] * gd[1] + g.d[2] * gd[2];
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float gu = (gd[a] - g.d[a] * dot) * g.inv_norm;
            acc[3 * a] = fmaf(gu, dir[0], acc[3 * a]);
            acc[3 * a + 1] = fmaf(gu, dir[1], acc[3 * a + 1]);
            acc[3 * a + 2] = fmaf(gu, dir[2], acc[3 * a + 2]);
            acc[9 + a] += go[a];
        }
    }
#pragma unroll
    for (int k = 0; k < 12; ++k) acc[k] = warp_sum(acc[k]);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 12; ++k) red[warp][k] = acc[k];
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        float t = 0.0f;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) t += red[w][threadIdx.x];
        atomicAdd(threadIdx.x < 9 ? d_rot + threadIdx.x : d_origin + (threadIdx.x - 9), t);
    }
}

// Camera rays (reference CameraRayDirections.build_rays, src/common/ray_utils.py:168-206, called through
// KeyFrame.build_camera_rays, src/mapping/keyframe.py:108-128, with the pixel selection of
// src/mapping/optimizer.py:626-651 fused in): directions [n,3] per pixel (camera frame), pixel grid coordinates,
// optional image [n,C] to gather the target intensities from.  far = cube exit only (no range clamp); the camera
// pose is detached in the reference (detach_rgb_from_poses), so there is no backward.
//   rays[i] = [o, d, -d, pix_i, pix_j, near, exit(o, d)];  intensities[i] = image[idx_i]
__global__ void __launch_bounds__(kThreads)
build_camera_rays_kernel(const float* __restrict__ directions, const float* __restrict__ pix_i,
                         const float* __restrict__ pix_j, const float* __restrict__ image, int n_colors,
                         const long long* __restrict__ indices, int n, int m, const float* __restrict__ rot,
                         const float* __restrict__ origin, float near, float* __restrict__ rays,
                         float* __restrict__ intensities) {
    const int i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= m) return;
    const Pose p = load_pose(rot, origin);
    long long j = indices ? indices[i] : (long long)i;
    j = j < 0 ? 0 : (j >= n ? (long long)n - 1 : j);       // never read out of bounds (the host wrapper validates)
    if (j < 0) j += n;
    const float* dir = directions + 3 * (size_t)j;
    const RayGeom g = ray_geometry(p, __ldg(dir), __ldg(dir + 1), __ldg(dir + 2));
    float* r = rays + (size_t)i * LB_RAY_STRIDE;
    r[0] = p.o[0]; r[1] = p.o[1]; r[2] = p.o[2];
    r[3] = g.d[0]; r[4] = g.d[1]; r[5] = g.d[2];
    r[6] = -g.d[0]; r[7] = -g.d[1]; r[8] = -g.d[2];
    r[9] = pix_i ? __ldg(pix_i + j) : 0.0f;
    r[10] = pix_j ? __ldg(pix_j + j) : 0.0f;
    r[11] = near;
    r[12] = g.exit_t;
    if (image)
        for (int c = 0; c < n_colors; ++c) intensities[(size_t)i * n_colors + c] = __ldg(image + (size_t)j * n_colors + c);
}

// ------------------------------------------------------------------------------------------
// pose 6-vector [t, axis-angle] -> (R [3,3] row-major, ray origin o = (t + shift) / scale), forward and backward, one
// thread per keyframe.  Same arithmetic as reference pose_utils.py:354-368 (pytorch3d axis_angle_to_matrix: axis-angle ->
// quaternion -> matrix, sin(x/2)/x replaced by its Taylor series below 1e-6) and ray_utils.py:293-296; the backward is
// the chain rule through exactly those formulas, so it equals what autograd produces for them.  In PyTorch this chain
// is ~40 tiny launches forward and ~250 backward per iteration of pose refinement; here it is two.
// ------------------------------------------------------------------------------------------
struct Quat {
    float th, sn, r, i, j, k, ts;     // angle, sin(th/2)/th, quaternion (r, i, j, k), 2 / |q|^2
};

__device__ __forceinline__ Quat pose_quat(const float* a) {
    Quat q;
    q.th = sqrtf(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
    const float h = 0.5f * q.th;
    q.sn = (fabsf(q.th) < 1e-6f) ? 0.5f - q.th * q.th / 48.0f : sinf(h) / q.th;
    q.r = cosf(h);
    q.i = a[0] * q.sn; q.j = a[1] * q.sn; q.k = a[2] * q.sn;
    q.ts = 2.0f / (q.r * q.r + q.i * q.i + q.j * q.j + q.k * q.k);
    return q;
}

__global__ void pose_to_rot_origin_kernel(const float* __restrict__ pose6, const float* __restrict__ shift, float scale,
                                          int K, float* __restrict__ rot, float* __restrict__ origin) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= K) return;
    const float* p = pose6 + 6 * n;
    const Quat q = pose_quat(p + 3);
    const float r = q.r, i = q.i, j = q.j, k = q.k, ts = q.ts;
    float* R = rot + 9 * n;
    R[0] = 1.0f - ts * (j * j + k * k); R[1] = ts * (i * j - k * r);        R[2] = ts * (i * k + j * r);
    R[3] = ts * (i * j + k * r);        R[4] = 1.0f - ts * (i * i + k * k); R[5] = ts * (j * k - i * r);
    R[6] = ts * (i * k - j * r);        R[7] = ts * (j * k + i * r);        R[8] = 1.0f - ts * (i * i + j * j);
#pragma unroll
    for (int c = 0; c < 3; ++c) origin[3 * n + c] = __fdiv_rn(__fadd_rn(p[c], __ldg(shift + c)), scale);
}

__global__ void pose_to_rot_origin_bwd_kernel(const float* __restrict__ pose6, float scale, int K,
                                              const float* __restrict__ d_rot, const float* __restrict__ d_origin,
                                              float* __restrict__ d_pose6) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= K) return;
    const float* a = pose6 + 6 * n + 3;
    const Quat q = pose_quat(a);
    const float r = q.r, i = q.i, j = q.j, k = q.k, ts = q.ts;
    const float* G = d_rot + 9 * n;
    // R = I + ts * M(q):  dL/dts = <G, M>,  dL/dq = ts * dM^T G  (+ dL/dts * dts/dq, dts/dq = -ts^2 q)
    const float m0 = -(j * j + k * k), m1 = i * j - k * r, m2 = i * k + j * r, m3 = i * j + k * r, m4 = -(i * i + k * k),
                m5 = j * k - i * r, m6 = i * k - j * r, m7 = j * k + i * r, m8 = -(i * i + j * j);
    const float dts = G[0] * m0 + G[1] * m1 + G[2] * m2 + G[3] * m3 + G[4] * m4 + G[5] * m5 + G[6] * m6 + G[7] * m7 + G[8] * m8;
    float dr = ts * (-G[1] * k + G[2] * j + G[3] * k - G[5] * i - G[6] * j + G[7] * i);
    float di = ts * (G[1] * j + G[2] * k + G[3] * j - 2.0f * G[4] * i - G[5] * r + G[6] * k + G[7] * r - 2.0f * G[8] * i);
    float dj = ts * (-2.0f * G[0] * j + G[1] * i + G[2] * r + G[3] * i + G[5] * k - G[6] * r + G[7] * k - 2.0f * G[8] * j);
    float dk = ts * (-2.0f * G[0] * k - G[1] * r + G[2] * i + G[3] * r - 2.0f * G[4] * k + G[5] * j + G[6] * i + G[7] * j);
    const float tt = -ts * ts * dts;
    dr += tt * r; di += tt * i; dj += tt * j; dk += tt * k;
    // q = (cos(th/2), a * sn(th)):  d a_c = sn * d(ijk)_c + [ (a . d(ijk)) * sn'(th) - dr * sin(th/2) / 2 ] * a_c / th
    const float h = 0.5f * q.th;
    const float dsn = (fabsf(q.th) < 1e-6f) ? -q.th / 24.0f : (0.5f * cosf(h) * q.th - sinf(h)) / (q.th * q.th);
    const float radial = (a[0] * di + a[1] * dj + a[2] * dk) * dsn - 0.5f * sinf(h) * dr;
    const float inv_th = q.th > 0.0f ? 1.0f / q.th : 0.0f;        // d|a|/da at a = 0: the subgradient 0, like torch.norm
    float* out = d_pose6 + 6 * n;
    out[3] = q.sn * di + radial * a[0] * inv_th;
    out[4] = q.sn * dj + radial * a[1] * inv_th;
    out[5] = q.sn * dk + radial * a[2] * inv_th;
#pragma unroll
    for (int c = 0; c < 3; ++c) out[c] = d_origin[3 * n + c] / scale;
}

}  // namespace

extern "C" int lb_pose_to_rot_origin(const float* pose6, const float* shift, float scale, int n_poses, float* rot,
                                     float* origin, void* stream) {
    LB_REQUIRE(n_poses == 0 || (pose6 && shift && rot && origin), "null pointer");
    LB_REQUIRE(n_poses >= 0 && scale > 0.0f, "bad sizes");
    if (n_poses == 0) return LB_OK;
    pose_to_rot_origin_kernel<<<(n_poses + 63) / 64, 64, 0, (cudaStream_t)stream>>>(pose6, shift, scale, n_poses, rot, origin);
    LB_LAUNCH_CHECK("lb_pose_to_rot_origin");
    return LB_OK;
}

extern "C" int lb_pose_to_rot_origin_bwd(const float* pose6, float scale, int n_poses, const float* d_rot,
                                         const float* d_origin, float* d_pose6, void* stream) {
    LB_REQUIRE(n_poses == 0 || (pose6 && d_rot && d_origin && d_pose6), "null pointer");
    LB_REQUIRE(n_poses >= 0 && scale > 0.0f, "bad sizes");
    if (n_poses == 0) return LB_OK;
    pose_to_rot_origin_bwd_kernel<<<(n_poses + 63) / 64, 64, 0, (cudaStream_t)stream>>>(pose6, scale, n_poses, d_rot, d_origin, d_pose6);
    LB_LAUNCH_CHECK("lb_pose_to_rot_origin_bwd");
    return LB_OK;
}

extern "C" int lb_build_lidar_rays(const float* directions, const float* distances, const int64_t* indices, int n,
                                   int m, const float* rot, const float* origin, float near, float far_max,
                                   float inv_scale, float const_depth, float mask_span, float* rays, float* depths,
                                   void* stream) {
    LB_REQUIRE(m == 0 || (directions && rot && origin && rays && depths), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n > 0 && m >= 0, "bad sizes");
    if (m == 0) return LB_OK;
    build_rays_kernel<<<(m + kThreads - 1) / kThreads, kThreads, 0, (cudaStream_t)stream>>>(
        directions, distances, reinterpret_cast<const long long*>(indices), n, m, rot, origin, near, far_max, inv_scale,
        const_depth, mask_span, rays, depths);
    LB_LAUNCH_CHECK("lb_build_lidar_rays");
    return LB_OK;
}

extern "C" int lb_build_lidar_rays_bwd(const float* directions, const int64_t* indices, int n, int m,
                                       const float* rot, const float* origin, float far_max, const float* d_rays,
                                       float* d_rot, float* d_origin, void* stream) {
    LB_REQUIRE(m == 0 || (directions && rot && origin && d_rays && d_rot && d_origin), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n > 0 && m >= 0, "bad sizes");
    if (m == 0) return LB_OK;
    int blocks = (m + kThreads - 1) / kThreads;
    if (blocks > 148) blocks = 148;
    build_rays_bwd_kernel<<<blocks, kThreads, 0, (cudaStream_t)stream>>>(
        directions, reinterpret_cast<const long long*>(indices), n, m, rot, origin, far_max, d_rays, d_rot, d_origin);
    LB_LAUNCH_CHECK("lb_build_lidar_rays_bwd");
    return LB_OK;
}

extern "C" int lb_build_camera_rays(const float* directions, const float* pix_i, const float* pix_j, const float* image,
                                    int n_colors, const int64_t* indices, int n, int m, const float* rot,
                                    const float* origin, float near, float* rays, float* intensities, void* stream) {
    LB_REQUIRE(m == 0 || (directions && rot && origin && rays), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n > 0 && m >= 0, "bad sizes");
    LB_REQUIRE(!image || (intensities && n_colors >= 1), "image needs an intensities output and n_colors >= 1");
    if (m == 0) return LB_OK;
    build_camera_rays_kernel<<<(m + kThreads - 1) / kThreads, kThreads, 0, (cudaStream_t)stream>>>(
        directions, pix_i, pix_j, image, n_colors, reinterpret_cast<const long long*>(indices), n, m, rot, origin, near,
        rays, intensities);
    LB_LAUNCH_CHECK("lb_build_camera_rays");
    return LB_OK;
}
