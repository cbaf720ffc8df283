// f3: the mesher's max-weight voxel splat (reference analysis/mesher.py:137-180).
//
// For every sample point xyz = o + d * z of every kept ray: if it lies inside the meshing bound, bucketize each
// coordinate against the axis's grid coordinates (torch.bucketize, right=False: the number of boundaries strictly
// below the value, evaluated in fp64 like the reference's float64 boundaries) and keep the LARGEST compositing
// weight seen in that voxel.  The reference writes results[idx] = max(results[idx], w) with duplicate indices in
// one indexed assignment (last writer wins, order undefined); here it is an exact atomic max, so the volume is the
// deterministic upper envelope of what the reference can produce.  Volume layout: [y][x][z] (np.meshgrid 'xy').
#include "lb_common.cuh"

namespace {

__device__ __forceinline__ int lower_bound(const double* __restrict__ b, int n, double v) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(b + mid) < v) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(256)
voxel_splat_max_kernel(const float* __restrict__ rays, const float* __restrict__ z, const float* __restrict__ w,
                       const uint8_t* __restrict__ ray_mask, int n_rays, int n_samples,
                       const double* __restrict__ xb, int nx, const double* __restrict__ yb, int ny,
                       const double* __restrict__ zb, int nz, float* __restrict__ volume) {
    const size_t total = (size_t)n_rays * n_samples;
    const long long n_vox = (long long)nx * ny * nz;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
        const size_t ray = i / n_samples;
        if (ray_mask && !ray_mask[ray]) continue;
        const float* r = rays + ray * LB_RAY_STRIDE;
        const float zi = __ldg(z + i);
        // points_fine = rays_o + rays_d * z as two rounded fp32 operations (rendering_tcnn.py:390)
        const float px = __fadd_rn(__ldg(r), __fmul_rn(__ldg(r + 3), zi));
        const float py = __fadd_rn(__ldg(r + 1), __fmul_rn(__ldg(r + 4), zi));
        const float pz = __fadd_rn(__ldg(r + 2), __fmul_rn(__ldg(r + 5), zi));
        const double dx = px, dy = py, dz = pz;
        if (!(dx >= __ldg(xb) && dx <= __ldg(xb + nx - 1) && dy >= __ldg(yb) && dy <= __ldg(yb + ny - 1) &&
              dz >= __ldg(zb) && dz <= __ldg(zb + nz - 1)))
            continue;
        const long long idx = (long long)lower_bound(xb, nx, dx) * nz + (long long)lower_bound(yb, ny, dy) * nx * nz +
                              lower_bound(zb, nz, dz);
        if (idx >= n_vox) continue;                                   // "hack around bucketize edge cases" (mesher.py:172)
        const float wv = __ldg(w + i);
        if (wv > 0.0f) atomicMax(reinterpret_cast<int*>(volume + idx), __float_as_int(wv));   // w >= 0: int order = float order
    }
}

}  // namespace

extern "C" int lb_voxel_splat_max(const float* rays, const float* z, const float* weights, const uint8_t* ray_mask,
                                  int n_rays, int n_samples, const double* x_bounds, int nx, const double* y_bounds,
                                  int ny, const double* z_bounds, int nz, float* volume, void* stream) {
    LB_REQUIRE(n_rays == 0 || (rays && z && weights && x_bounds && y_bounds && z_bounds && volume), "null pointer");   // an empty batch may pass NULL buffers
    LB_REQUIRE(n_rays >= 0 && n_samples >= 1 && nx >= 1 && ny >= 1 && nz >= 1, "bad sizes");
    if (n_rays == 0) return LB_OK;
    const size_t total = (size_t)n_rays * n_samples;
    size_t blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    voxel_splat_max_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(rays, z, weights, ray_mask, n_rays, n_samples,
                                                                              x_bounds, nx, y_bounds, ny, z_bounds, nz, volume);
    LB_LAUNCH_CHECK("lb_voxel_splat_max");
    return LB_OK;
}
