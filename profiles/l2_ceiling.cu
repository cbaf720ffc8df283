// L2 request-rate ceiling for the access patterns of the two fused sigma kernels (profiles/README.md, round 2).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/l2_ceiling profiles/l2_ceiling.cu
//   ./profiles/l2_ceiling [table_MB=57] [json_out]
//
// No math, no tensor cores, no shared memory: every thread issues randomly addressed accesses into one table that
// is L2-resident (57 MB = the T = 2^19 hash tables, or their gradient), with U independent accesses in flight per
// thread and full occupancy.  Patterns (one "request" = one 32-byte L2 sector touched by one warp instruction):
//   ld8       ld.global.nc.v2.f32        one random 8-byte entry per lane                 32 sectors / instruction
//   ld8pair   ld.global.nc.v2.f32        lanes 2k, 2k+1 read entries idx, idx^1           16 sectors / instruction
//                                        (the two x-neighbours of the encode kernels when the cell's x is even)
//   ld16      ld.global.nc.v4.f32        one random 16-byte aligned pair per lane          32 sectors / instruction
//   ld8mix    like the encode kernels: lane pairs read (idx, idx^1) for half of the points and two unrelated
//             entries for the other half (cell x odd)                                      24 sectors / instruction
//   red8, red8pair, red16 (red.global.add.v4.f32), red8mix: the same with red.global.add reductions
// Output: G sectors/s, G warp-instructions/s and GB/s (sectors x 32 B) per pattern, as text and as JSON.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CHECK(x)                                                                                       \
    do {                                                                                               \
        cudaError_t e_ = (x);                                                                          \
        if (e_ != cudaSuccess) {                                                                       \
            fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_));         \
            exit(1);                                                                                   \
        }                                                                                              \
    } while (0)

enum Pattern { LD8 = 0, LD8PAIR, LD16, LD8MIX, RED8, RED8PAIR, RED16, RED8MIX, FWDMIX, BWDMIX, N_PATTERNS };
static const char* kNames[N_PATTERNS] = {"ld8", "ld8pair", "ld16", "ld8mix", "red8", "red8pair", "red16", "red8mix",
                                         "fwdmix", "bwdmix"};
// sectors the SMs send to L2 per warp instruction of the random-access part (the unit of lts__t_sectors_srcunit_tex)
static const double kSectors[N_PATTERNS] = {32, 16, 32, 24, 32, 16, 32, 24, 24, 24};
// composite patterns (round 2): the request mix of the two fused sigma kernels, per iteration of 8 random accesses
//   fwdmix  8 x ld8mix (192 sectors per warp) + 4 coalesced streaming 8-byte stores per lane (32 sectors): 14 % writes,
//           sigma_fwd_fused_kernel's ratio (354 M table-read sectors : 69 M sectors of saved encodings per launch)
//   bwdmix  8 x red8mix (192 sectors) + 2 coalesced streaming 8-byte loads per lane (16 sectors): 8 % reads,
//           sigma_bwd_fused_kernel's ratio (444 M reduction sectors : ~40 M sectors of encodings / masks / d_sigma)
static const double kExtraSectors[N_PATTERNS] = {0, 0, 0, 0, 0, 0, 0, 0, 32, 16};

__device__ __forceinline__ uint32_t lcg(uint32_t& s) {
    s = s * 1664525u + 1013904223u;
    return s;
}
__device__ __forceinline__ float2 ldg8(const float2* p) {
    float2 v;
    asm volatile("ld.global.nc.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 ldg16(const float4* p) {
    float4 v;
    asm volatile("ld.global.nc.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void red8(float* p, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void red16(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// index of this lane's access in iteration state s.  n_entries = float2 entries in the table (even).
template <int kPattern>
__device__ __forceinline__ uint32_t pick(uint32_t& s, uint32_t n_entries, int lane) {
    if (kPattern == LD8 || kPattern == RED8) return __umulhi(lcg(s), n_entries);
    if (kPattern == LD16 || kPattern == RED16) return __umulhi(lcg(s), n_entries >> 1) << 1;
    // pair patterns: both lanes of a pair draw the same number (the even lane's stream)
    const uint32_t r = lcg(s);
    const uint32_t base = __umulhi(r, n_entries >> 1) << 1;
    if (kPattern == LD8PAIR || kPattern == RED8PAIR) return base + (lane & 1);
    // (LD8MIX, RED8MIX, FWDMIX, BWDMIX share the mixed pairing below)
    // mix: even-x cells (r bit 0 clear) are adjacent entries, odd-x cells two unrelated entries
    if (r & 1u) return (lane & 1) ? __umulhi(r * 2654435761u, n_entries) : base;
    return base + (lane & 1);
}

template <int kPattern, int kUnroll>
__global__ void __launch_bounds__(256) l2_kernel(float2* table, uint32_t n_entries, int iters, float* sink, float2* stream,
                                                 size_t stream_entries) {
    const int lane = threadIdx.x & 31;
    const uint32_t gtid = blockIdx.x * blockDim.x + threadIdx.x;
    // pair patterns: lanes 2k and 2k+1 share a stream
    const bool paired = kPattern == LD8PAIR || kPattern == RED8PAIR || kPattern == LD8MIX || kPattern == RED8MIX ||
                        kPattern == FWDMIX || kPattern == BWDMIX;
    const size_t n_threads = (size_t)gridDim.x * blockDim.x;
    uint32_t s = (paired ? (gtid >> 1) : gtid) * 2654435761u + 12345u;
    float acc = 0.0f;
    for (int it = 0; it < iters; ++it) {
        uint32_t idx[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) idx[u] = pick<kPattern>(s, n_entries, lane);
        if (kPattern == FWDMIX) {            // streaming stores of use-once data next to the gathers
#pragma unroll
            for (int w = 0; w < 4; ++w) __stcs(stream + ((size_t)(it * 4 + w) * n_threads + gtid) % stream_entries, make_float2(acc, 1.0f));
        }
        if (kPattern == BWDMIX) {            // streaming loads of use-once data next to the reductions
#pragma unroll
            for (int w = 0; w < 2; ++w) {
                const float2 t = __ldcs(stream + ((size_t)(it * 2 + w) * n_threads + gtid) % stream_entries);
                acc += t.x;
            }
        }
        if (kPattern == LD8 || kPattern == LD8PAIR || kPattern == LD8MIX || kPattern == FWDMIX) {
            float2 v[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) v[u] = ldg8(table + idx[u]);
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) acc += v[u].x + v[u].y;
        } else if (kPattern == LD16) {
            float4 v[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) v[u] = ldg16(reinterpret_cast<const float4*>(table + idx[u]));
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) acc += v[u].x + v[u].w;
        } else if (kPattern == RED16) {
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) red16(reinterpret_cast<float*>(table + idx[u]), 1.0f, 1.0f, 1.0f, 1.0f);
        } else {
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) red8(reinterpret_cast<float*>(table + idx[u]), 1.0f, 1.0f);
        }
    }
    if (acc == 123.456f) sink[0] = acc;      // keeps the loads alive
}

template <int kPattern>
static double run(float2* table, uint32_t n_entries, float* sink, int blocks, int iters, double* ms_out, float2* stream,
                  size_t stream_entries) {
    constexpr int kUnroll = 8;
    cudaEvent_t e0, e1;
    CHECK(cudaEventCreate(&e0));
    CHECK(cudaEventCreate(&e1));
    for (int w = 0; w < 3; ++w) l2_kernel<kPattern, kUnroll><<<blocks, 256>>>(table, n_entries, iters / 4 + 1, sink, stream, stream_entries);
    CHECK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CHECK(cudaEventRecord(e0));
        l2_kernel<kPattern, kUnroll><<<blocks, 256>>>(table, n_entries, iters, sink, stream, stream_entries);
        CHECK(cudaEventRecord(e1));
        CHECK(cudaEventSynchronize(e1));
        float ms;
        CHECK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    CHECK(cudaGetLastError());
    *ms_out = best;
    const double warp_instr = (double)blocks * 8 * iters * kUnroll;
    return warp_instr / (best * 1e-3);       // warp instructions per second
}

int main(int argc, char** argv) {
    const double table_mb = argc > 1 ? atof(argv[1]) : 56.9;
    const char* json_path = argc > 2 ? argv[2] : nullptr;
    cudaDeviceProp prop;
    CHECK(cudaGetDeviceProperties(&prop, 0));
    const uint32_t n_entries = ((uint32_t)(table_mb * 1e6 / 8)) & ~1u;
    float2* table;
    float* sink;
    CHECK(cudaMalloc(&table, (size_t)n_entries * 8));
    CHECK(cudaMemset(table, 0, (size_t)n_entries * 8));
    CHECK(cudaMalloc(&sink, 4));
    const size_t stream_entries = (size_t)1 << 27;            // 1 GiB of float2: use-once streams never hit the L2
    float2* stream;
    CHECK(cudaMalloc(&stream, stream_entries * 8));
    CHECK(cudaMemset(stream, 0, stream_entries * 8));
    const int blocks = prop.multiProcessorCount * 8;          // 2048 threads per SM
    const int iters = 400;                                    // 148*8*8 warps x 400 x 8 = 30 M warp instr = 0.97 G lane accesses
    printf("device %s, %d SMs, L2 %d MB, table %.1f MB (%u entries), %d blocks x 256 threads, %d accesses in flight per thread\n",
           prop.name, prop.multiProcessorCount, prop.l2CacheSize >> 20, n_entries * 8 / 1e6, n_entries, blocks, 8);
    double wips[N_PATTERNS], ms[N_PATTERNS];
    wips[LD8] = run<LD8>(table, n_entries, sink, blocks, iters, &ms[LD8], stream, stream_entries);
    wips[LD8PAIR] = run<LD8PAIR>(table, n_entries, sink, blocks, iters, &ms[LD8PAIR], stream, stream_entries);
    wips[LD16] = run<LD16>(table, n_entries, sink, blocks, iters, &ms[LD16], stream, stream_entries);
    wips[LD8MIX] = run<LD8MIX>(table, n_entries, sink, blocks, iters, &ms[LD8MIX], stream, stream_entries);
    wips[RED8] = run<RED8>(table, n_entries, sink, blocks, iters, &ms[RED8], stream, stream_entries);
    wips[RED8PAIR] = run<RED8PAIR>(table, n_entries, sink, blocks, iters, &ms[RED8PAIR], stream, stream_entries);
    wips[RED16] = run<RED16>(table, n_entries, sink, blocks, iters, &ms[RED16], stream, stream_entries);
    wips[RED8MIX] = run<RED8MIX>(table, n_entries, sink, blocks, iters, &ms[RED8MIX], stream, stream_entries);
    wips[FWDMIX] = run<FWDMIX>(table, n_entries, sink, blocks, iters, &ms[FWDMIX], stream, stream_entries);
    wips[BWDMIX] = run<BWDMIX>(table, n_entries, sink, blocks, iters, &ms[BWDMIX], stream, stream_entries);
    FILE* jf = json_path ? fopen(json_path, "w") : nullptr;
    if (jf) fprintf(jf, "{\"device\": \"%s\", \"table_mb\": %.1f, \"threads_per_sm\": 2048, \"in_flight_per_thread\": 8, \"patterns\": {", prop.name, n_entries * 8 / 1e6);
    for (int p = 0; p < N_PATTERNS; ++p) {
        const double sect = wips[p] * (kSectors[p] + kExtraSectors[p] / 8.0);       // the extra stream accesses come once per 8 random ones
        printf("%-9s %8.3f ms  %7.2f G warp-instr/s  %7.1f G sectors/s  %8.1f GB/s of sectors  %7.1f G lane-accesses/s\n",
               kNames[p], ms[p], wips[p] / 1e9, sect / 1e9, sect * 32 / 1e9, wips[p] * 32 / 1e9);
        if (jf) fprintf(jf, "%s\"%s\": {\"ms\": %.4f, \"g_warp_instr_per_s\": %.3f, \"g_sectors_per_s\": %.2f, \"sector_gbs\": %.1f, \"g_lane_accesses_per_s\": %.2f}",
                        p ? ", " : "", kNames[p], ms[p], wips[p] / 1e9, sect / 1e9, sect * 32 / 1e9, wips[p] * 32 / 1e9);
    }
    if (jf) {
        fprintf(jf, "}}\n");
        fclose(jf);
    }
    return 0;
}
