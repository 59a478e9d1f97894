// Streaming ceilings on this B200: what a read-only pass and a read+write pass over a tensor
// larger than L2 can reach with (a) plain vector loads and (b) the TMA pipelines used by
// libk7det.  Build + run (GPU box):
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o /tmp/membench tools/membench.cu && /tmp/membench
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../objectdetection.pytorch_b200/csrc/k7_common.cuh"
void k7_set_error(const char*, ...) {}

#define CK(x)                                                                      \
    do {                                                                           \
        cudaError_t e = (x);                                                       \
        if (e != cudaSuccess) {                                                    \
            printf("%s: %s\n", #x, cudaGetErrorString(e));                         \
            exit(1);                                                               \
        }                                                                          \
    } while (0)

// (a) grid-stride LDG.128 read (sum) / copy
template <int UNROLL, bool WRITE>
__global__ void __launch_bounds__(512) ldg_kernel(const float4* __restrict__ in, float4* __restrict__ out, size_t n4,
                                                  float* __restrict__ sink) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    float acc = 0.f;
    for (; i + (UNROLL - 1) * stride < n4; i += UNROLL * stride) {
        float4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = __ldcs(in + i + u * stride);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            if (WRITE) __stcs(out + i + u * stride, v[u]);
            else acc += v[u].x + v[u].y + v[u].z + v[u].w;
        }
    }
    for (; i < n4; i += stride) {
        float4 v = in[i];
        if (WRITE) out[i] = v;
        else acc += v.x + v.y + v.z + v.w;
    }
    if (!WRITE && acc == 123.456f) *sink = acc;
}

// (b) CTA pipeline: producer warp + CW consumer warps, stages of CW*512 floats (the k7_loss layout).
// Consumers do `WORK` dummy FMAs per float4 so the issue pressure can be dialled.
template <bool WRITE, int WORK>
__global__ void __launch_bounds__(17 * 32, 1) pipe_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                          uint32_t num_tiles, int NS, int CW, float* __restrict__ sink) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t tile_f = CW * 512u, stride = tile_f * 4u;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)NS * stride);
    uint64_t* done = full + NS;
    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; ++s) {
            k7_mbar_init(full + s, 1);
            k7_mbar_init(done + s, CW);
        }
        k7_fence_mbar_init();
        k7_fence_proxy_async();
    }
    __syncthreads();
    if (blockIdx.x >= num_tiles) return;
    const uint32_t ntl = (num_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;
    if (warp == CW) {
        if (lane) return;
        auto load = [&](uint32_t j) {
            const size_t g = blockIdx.x + (size_t)j * gridDim.x;
            const int s = j % NS;
            k7_mbar_expect_tx(full + s, stride);
            k7_bulk_g2s(smem + (size_t)s * stride, in + g * tile_f, stride, full + s);
        };
        const uint32_t ahead = WRITE ? NS - 1 : NS;
        uint32_t jl = 0;
        for (; jl < ntl && jl < ahead; ++jl) load(jl);
        for (uint32_t j = 0; j < ntl; ++j) {
            const int s = j % NS;
            k7_mbar_wait(done + s, (j / NS) & 1u);
            if (WRITE) {
                const size_t g = blockIdx.x + (size_t)j * gridDim.x;
                k7_bulk_s2g(out + g * tile_f, smem + (size_t)s * stride, stride);
                k7_bulk_commit();
                if (jl < ntl) {
                    k7_bulk_wait_read<1>();
                    load(jl++);
                }
            } else if (jl < ntl) {
                load(jl++);
            }
        }
        if (WRITE) k7_bulk_wait<0>();
        return;
    }
    float acc = 0.f;
    for (uint32_t j = 0; j < ntl; ++j) {
        const int s = j % NS;
        float4* t4 = reinterpret_cast<float4*>(smem + (size_t)s * stride) + warp * 128;
        k7_mbar_wait(full + s, (j / NS) & 1u);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float4 v = t4[lane + 32 * k];
#pragma unroll
            for (int w = 0; w < WORK; ++w) {
                v.x = fmaf(v.x, 1.0001f, 0.5f); v.y = fmaf(v.y, 1.0001f, 0.5f);
                v.z = fmaf(v.z, 1.0001f, 0.5f); v.w = fmaf(v.w, 1.0001f, 0.5f);
            }
            if (WRITE) t4[lane + 32 * k] = v;
            else acc += v.x + v.y + v.z + v.w;
        }
        if (WRITE) k7_fence_proxy_async();
        __syncwarp();
        if (lane == 0) k7_mbar_arrive(done + s);
    }
    if (!WRITE && acc == 123.456f) *sink = acc;
}

template <typename F>
static float timeit(F f, int iters = 10) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    for (int i = 0; i < 3; ++i) f();
    CK(cudaDeviceSynchronize());
    cudaEventRecord(e0);
    for (int i = 0; i < iters; ++i) f();
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    return ms / iters;
}

int main() {
    const size_t n = (size_t)64 * 49104 * 90 / 8192 * 8192;   // ~1.13 GB, multiple of the tile
    float *in, *out, *sink;
    CK(cudaMalloc(&in, n * 4));
    CK(cudaMalloc(&out, n * 4));
    CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(in, 0, n * 4));
    const double gb = n * 4 / 1e9;
    int sms;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    printf("tensor %.3f GB, %d SMs\n", gb, sms);
    float ms = timeit([&] { CK(cudaMemcpyAsync(out, in, n * 4, cudaMemcpyDeviceToDevice)); });
    printf("cudaMemcpy D2D              : %7.1f us  %6.0f GB/s (r+w)\n", ms * 1e3, 2 * gb / ms * 1e3);
    for (int mult : {2, 4, 8}) {
        ms = timeit([&] { ldg_kernel<4, false><<<sms * mult, 512>>>((float4*)in, (float4*)out, n / 4, sink); });
        printf("LDG.128 x4 read  grid %2dxSM : %7.1f us  %6.0f GB/s (r)\n", mult, ms * 1e3, gb / ms * 1e3);
        ms = timeit([&] { ldg_kernel<4, true><<<sms * mult, 512>>>((float4*)in, (float4*)out, n / 4, sink); });
        printf("LDG.128 x4 copy  grid %2dxSM : %7.1f us  %6.0f GB/s (r+w)\n", mult, ms * 1e3, 2 * gb / ms * 1e3);
    }
    ms = timeit([&] { ldg_kernel<8, false><<<sms * 4, 512>>>((float4*)in, (float4*)out, n / 4, sink); });
    printf("LDG.128 x8 read  grid  4xSM : %7.1f us  %6.0f GB/s (r)\n", ms * 1e3, gb / ms * 1e3);

    const int CW = 16;
    const uint32_t tiles = (uint32_t)(n / (CW * 512));
    for (int NS : {3, 4, 6}) {
        const size_t smem = (size_t)NS * CW * 2048 + NS * 16;
#define RUN(WR, WK)                                                                                              \
    do {                                                                                                         \
        auto k = pipe_kernel<WR, WK>;                                                                            \
        CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                     \
        ms = timeit([&] { k<<<sms, (CW + 1) * 32, smem>>>(in, out, tiles, NS, CW, sink); });                     \
        CK(cudaGetLastError());                                                                                  \
        printf("TMA pipe %s NS=%d work=%2d     : %7.1f us  %6.0f GB/s\n", WR ? "copy" : "read", NS, WK, ms * 1e3, \
               (WR ? 2 : 1) * gb / ms * 1e3);                                                                    \
    } while (0)
        RUN(false, 0);
        RUN(false, 8);
        RUN(false, 16);
        RUN(true, 0);
        RUN(true, 8);
        RUN(true, 16);
        RUN(true, 24);
    }
    return 0;
}
