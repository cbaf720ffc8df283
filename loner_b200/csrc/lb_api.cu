// Plan / error handling of the C-ABI (include/loner_b200.h).
#include <math.h>
#include <stdarg.h>
#include <stdlib.h>

#include "lb_common.cuh"

static thread_local char g_err[512] = "";

void lb_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int lb_check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return LB_OK;
    lb_set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
    return LB_ERR_CUDA;
}

extern "C" int lb_version(void) { return LB_VERSION; }

extern "C" const char* lb_last_error(void) { return g_err; }

// Level table following tiny-cuda-nn's GridEncodingTemplated constructor and grid_index():
// scale_l = exp2(l*log2(pls))*base - 1; res = ceil(scale)+1; params = min(roundup8(res^3), 2^log2T);
// dense iff the running stride never exceeds the level size.
extern "C" int lb_plan_create(const lb_grid_config* cfg, lb_plan** out) {
    LB_REQUIRE(cfg && out, "null argument");
    if (cfg->n_features_per_level != LB_FEATS || cfg->mlp_width != LB_MLP_W || cfg->n_levels < 1 ||
        cfg->n_levels > 16 || cfg->mlp_hidden_layers != 1) {
        lb_set_error("lb_plan_create: unsupported config (need F=2, width=64, 1 hidden layer, <=16 levels)");
        return LB_ERR_UNSUPPORTED;
    }
    LB_REQUIRE(cfg->log2_hashmap_size >= 4 && cfg->log2_hashmap_size <= 28, "log2_hashmap_size out of range");
    LB_REQUIRE(cfg->base_resolution >= 1 && cfg->per_level_scale >= 1.0f, "bad resolution / scale");
    lb_plan* p = (lb_plan*)calloc(1, sizeof(lb_plan));
    if (!p) {
        lb_set_error("lb_plan_create: out of host memory");
        return LB_ERR_BAD_ARGUMENT;
    }
    p->cfg = *cfg;
    const float log2_pls = log2f(cfg->per_level_scale);
    uint64_t offset = 0;
    for (int l = 0; l < cfg->n_levels; ++l) {
        const float scale = exp2f((float)l * log2_pls) * (float)cfg->base_resolution - 1.0f;
        const uint32_t res = (uint32_t)ceilf(scale) + 1u;
        const uint32_t max_params = 0xFFFFFFFFu / 2u;
        uint64_t dense_params = (uint64_t)res * res * res;
        uint32_t params = (powf((float)res, 3.0f) > (float)max_params) ? max_params : (uint32_t)dense_params;
        params = (params + 7u) / 8u * 8u;
        const uint32_t T = 1u << cfg->log2_hashmap_size;
        if (params > T) params = T;
        uint64_t stride = 1;
        for (int d = 0; d < 3 && stride <= params; ++d) stride *= res;
        p->levels[l].scale = scale;
        p->levels[l].res = res;
        p->levels[l].res2 = res * res;
        p->levels[l].offset = (uint32_t)offset;
        p->levels[l].size = params;
        p->levels[l].dense = (params < stride) ? 0u : 1u;
        p->levels[l].mask = ((params & (params - 1u)) == 0u) ? params - 1u : 0u;
        offset += params;
    }
    if (offset * LB_FEATS > 0x7FFFFFFFull) {
        free(p);
        lb_set_error("lb_plan_create: table too large for 32-bit indexing");
        return LB_ERR_UNSUPPORTED;
    }
    p->n_table_entries = (int64_t)offset;
    p->n_mlp_params = LB_MLP_W * (cfg->n_levels * LB_FEATS) + LB_MLP_OUT_PAD * LB_MLP_W;
    int dev = 0;
    p->sm_count = 148;
    p->l2_bytes = 126ll << 20;
    if (cudaGetDevice(&dev) == cudaSuccess) {
        int sm = 0, l2 = 0;
        if (cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && sm > 0) p->sm_count = sm;
        if (cudaDeviceGetAttribute(&l2, cudaDevAttrL2CacheSize, dev) == cudaSuccess && l2 > 0) p->l2_bytes = l2;
    }
    cudaGetLastError();  // plan creation must work without a device (host-only tests)
    *out = p;
    return LB_OK;
}

extern "C" void lb_plan_destroy(lb_plan* plan) { free(plan); }

extern "C" int lb_plan_set_l2_bytes(lb_plan* plan, int64_t l2_bytes) {
    LB_REQUIRE(plan && l2_bytes > 0, "null plan or non-positive size");
    plan->l2_bytes = l2_bytes;
    return LB_OK;
}

extern "C" int lb_build_flags(void) {
#ifdef LB_TEST_SWITCHES
    return 1;
#else
    return 0;
#endif
}

extern "C" int lb_plan_param_counts(const lb_plan* plan, int64_t* n_mlp, int64_t* n_table) {
    LB_REQUIRE(plan, "null plan");
    if (n_mlp) *n_mlp = plan->n_mlp_params;
    if (n_table) *n_table = plan->n_table_entries * LB_FEATS;
    return LB_OK;
}

extern "C" int lb_plan_level(const lb_plan* plan, int level, float* scale, uint32_t* resolution,
                             uint32_t* offset, uint32_t* size, int32_t* is_dense) {
    LB_REQUIRE(plan, "null plan");
    LB_REQUIRE(level >= 0 && level < plan->cfg.n_levels, "level out of range");
    const lb_level& L = plan->levels[level];
    if (scale) *scale = L.scale;
    if (resolution) *resolution = L.res;
    if (offset) *offset = L.offset;
    if (size) *size = L.size;
    if (is_dense) *is_dense = (int32_t)L.dense;
    return LB_OK;
}
