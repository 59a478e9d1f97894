// Shared host/device plumbing for libk7det.so (sm_100a only).
#pragma once
#include <cstdlib>

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/k7det.h"
#include "k7_math.cuh"

// ------------------------------------------------------------------ errors
void k7_set_error(const char* fmt, ...);

#define K7_REQUIRE(cond, ...)                \
    do {                                     \
        if (!(cond)) {                       \
            k7_set_error(__VA_ARGS__);       \
            return K7_EINVAL;                \
        }                                    \
    } while (0)

#define K7_CUDA(call)                                                                  \
    do {                                                                               \
        cudaError_t e_ = (call);                                                       \
        if (e_ != cudaSuccess) {                                                       \
            k7_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_),       \
                         __FILE__, __LINE__);                                          \
            return K7_ECUDA;                                                           \
        }                                                                              \
    } while (0)

#define K7_LAUNCH_CHECK()              \
    do {                               \
        k7_count_launch();             \
        K7_CUDA(cudaGetLastError());   \
    } while (0)

int k7_check_spec(const K7AnchorSpec* spec, int A);
int k7_sm_count();

static inline size_t k7_align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

void k7_count_launch();   // every kernel launch of the library ticks k7_launch_count()

// ------------------------------------------------------------------ programmatic dependent launch
// Chains of short dependent kernels (sort -> mask -> resolve) pay a launch gap per link.  A kernel
// launched with k7_launch_pdl may be scheduled while its predecessor in the stream is still running
// (once every CTA of the predecessor has called k7_pdl_trigger() or exited); it must call
// k7_pdl_wait() before it touches anything the predecessor wrote - the wait returns when the
// predecessor grid has completed and its memory is visible.  Both calls are no-ops in a normal launch.
#ifdef __CUDACC__
__device__ __forceinline__ void k7_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void k7_pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// experiment switch: K7_PDL=<bitmask>, bit i enables the attribute for the i-th k7_launch_pdl call site class
static inline int& k7_pdl_site() { static thread_local int site = 0; return site; }
static inline bool k7_pdl_mask_next() {
    static const int mask = [] { const char* e = getenv("K7_PDL"); return e ? atoi(e) : 14;   /* measured: not for the sort kernel */ }();
    return (mask >> k7_pdl_site()) & 1;
}
template <typename... KArgs, typename... Args>
static inline cudaError_t k7_launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                        Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = k7_pdl_mask_next() ? 1 : 0;
    k7_count_launch();
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#endif

// ------------------------------------------------------------------ division by a runtime constant
struct K7FastDiv {
    uint32_t m, sh;
};
// exact x / d for x < 2^31 (Granlund-Montgomery, round-up multiplier)
static inline K7FastDiv k7_fastdiv_make(uint32_t d) {
    int s = 0;
    while ((1ull << s) < d) ++s;
    K7FastDiv f;
    f.m = (uint32_t)(((1ull << (31 + s)) + d - 1) / d);
    f.sh = (uint32_t)(31 + s);
    return f;
}
__device__ __forceinline__ uint32_t k7_div(uint32_t x, const K7FastDiv f) {
    return (uint32_t)(((uint64_t)x * f.m) >> f.sh);
}

// ------------------------------------------------------------------ anchors on the fly
// Anchor `a` of the recipe, returned as (x1,y1,x2,y2) whatever the recipe's storage order.
// float64 arithmetic then ONE rounding to float32, exactly like
// architectures/anchors.py:99-119 + :58 (numpy float64 -> astype(float32)).
__device__ __forceinline__ void k7_anchor_from_spec(const K7AnchorSpec& s, int a, float& x1, float& y1,
                                                    float& x2, float& y2) {
    int l = 0;
#pragma unroll
    for (int i = 1; i < K7_MAX_LEVELS; ++i)
        if (i < s.num_levels && a >= s.offset[i]) l = i;
    const int r = a - s.offset[l];
    const int cell = r / s.cell_anchors;
    const int k = r - cell * s.cell_anchors;
    const int fw = s.fw[l];
    const int y = cell / fw;
    const int x = cell - y * fw;
    const double st = s.stride[l];
    const double sx = __dmul_rn(__dadd_rn((double)x, 0.5), st);
    const double sy = __dmul_rn(__dadd_rn((double)y, 0.5), st);
    const double* b = s.base[l][k];
    if (s.box_order == K7_BOX_XYXY) {
        x1 = __double2float_rn(__dadd_rn(b[0], sx));
        y1 = __double2float_rn(__dadd_rn(b[1], sy));
        x2 = __double2float_rn(__dadd_rn(b[2], sx));
        y2 = __double2float_rn(__dadd_rn(b[3], sy));
    } else {
        y1 = __double2float_rn(__dadd_rn(b[0], sy));
        x1 = __double2float_rn(__dadd_rn(b[1], sx));
        y2 = __double2float_rn(__dadd_rn(b[2], sy));
        x2 = __double2float_rn(__dadd_rn(b[3], sx));
    }
}

// Anchor from a materialised [A,4] tensor in `box_order` layout.
__device__ __forceinline__ void k7_anchor_from_tensor(const float* __restrict__ anchors, int box_order, int a,
                                                      float& x1, float& y1, float& x2, float& y2) {
    const float4 v = __ldg(reinterpret_cast<const float4*>(anchors) + a);
    if (box_order == K7_BOX_XYXY) {
        x1 = v.x; y1 = v.y; x2 = v.z; y2 = v.w;
    } else {
        y1 = v.x; x1 = v.y; y2 = v.z; x2 = v.w;
    }
}

template <bool FROM_SPEC>
__device__ __forceinline__ void k7_anchor(const K7AnchorSpec& s, const float* __restrict__ anchors, int box_order,
                                          int a, float& x1, float& y1, float& x2, float& y2) {
    if (FROM_SPEC)
        k7_anchor_from_spec(s, a, x1, y1, x2, y2);
    else
        k7_anchor_from_tensor(anchors, box_order, a, x1, y1, x2, y2);
}

// Layout of the caller's `best` buffer: [B*A] uint64 maxima, then (256-byte aligned) a header with one
// counter per image (K7_BEST_CNT_STRIDE words apart: atomics on ONE address serialise at ~7 ns each, measured;
// one counter per image keeps them parallel), then B lists of up to A anchors: the anchors of image b whose
// entry was written for the first time in this generation.  k7_loss walks the lists, not the array.
#define K7_BEST_CNT_STRIDE 16
__host__ __device__ inline size_t k7_best_list_off(int B, int A) {
    return ((sizeof(unsigned long long) * (size_t)B * (size_t)A + 255) / 256) * 256;
}
__host__ __device__ inline size_t k7_best_list_hdr(int B) {
    return ((sizeof(unsigned int) * K7_BEST_CNT_STRIDE * (size_t)B + 255) / 256) * 256;
}

// ------------------------------------------------------------------ warp helpers
__device__ __forceinline__ float k7_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ int k7_warp_sum(int v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------ mbarrier + bulk async copy (TMA 1-D)
// Blackwell/Hopper async-proxy primitives used by the streaming kernels.  SASS: UBLKCP / SYNCS.
__device__ __forceinline__ uint32_t k7_smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void k7_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(k7_smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void k7_fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void k7_fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void k7_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(k7_smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void k7_mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(k7_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void k7_mbar_wait_u32(uint32_t bar_smem_addr, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "K7_WAITA_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra K7_DONEA_%=;\n"
        "bra K7_WAITA_%=;\n"
        "K7_DONEA_%=:\n"
        "}\n" ::"r"(bar_smem_addr),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void k7_mbar_arrive_u32(uint32_t bar_smem_addr) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_smem_addr) : "memory");
}
__device__ __forceinline__ void k7_mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "K7_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra K7_DONE_%=;\n"
        "bra K7_WAIT_%=;\n"
        "K7_DONE_%=:\n"
        "}\n" ::"r"(k7_smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared, completion counted in bytes on `bar`.  16-byte aligned addresses, bytes % 16 == 0.
__device__ __forceinline__ void k7_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            k7_smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(k7_smem_u32(bar))
        : "memory");
}
// shared -> global, tracked by the issuing thread's bulk group.
__device__ __forceinline__ void k7_bulk_s2g(void* gmem_dst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst),
                 "r"(k7_smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void k7_bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void k7_bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void k7_bulk_wait() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
