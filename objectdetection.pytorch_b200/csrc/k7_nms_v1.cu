// The round-1 NMS chain, kept for batches of MANY images: one CTA per image ranks the survivors and lays them out
// in a spatial order (k7_sort_kernel), all SMs build a sparse suppression bitmask from 64-box blocks whose
// summaries can overlap (k7_mask_kernel), one CTA per image resolves greedy NMS as a fixed point
// (k7_resolve_kernel).  With >= K7_V1_MIN_IMAGES images in flight its per-image CTAs fill the GPU and it is the
// faster chain (C2: 66 us against 110 us for the multi-SM chain of k7_detect.cu, C3: 310 against 460); the
// multi-SM chain is for the few-image case (a shard of a large batch), where one CTA per image leaves most SMs
// idle.  Both produce torchvision.ops.nms's result exactly (tests run every NMS case through both).
//   replaces torchvision.ops.nms call sites     retinanet.py:314, utils/utils.py:93
#include <math.h>
#include <string.h>

#include "k7_common.cuh"
#include "k7_nms_v1.cuh"

namespace k7v1 {

#define K7_NMS_THREADS 1024
#define K7_NMS_CHUNK 1024
#define K7_NMS_WORDS (K7_NMS_CHUNK / 64)
#define K7_NMS_PITCH (K7_NMS_CHUNK + 1)   // u64 pitch of a mask word-row: conflict-free for writer and scanner

struct K7SortSmem {
    unsigned int hist[256];
    unsigned int base[256];
    unsigned int wtot[32];
    int skip;
    int pad[3];
    unsigned int warp_cnt[32][256];
};
struct K7NmsSmem {
    float4 box[K7_NMS_CHUNK];
    float area[K7_NMS_CHUNK];
    float4 kbox[256];
    float karea[256];
    int klabel[256];
    int label[K7_NMS_CHUNK];
    unsigned int alive_bits[K7_NMS_CHUNK / 32];
    int kept_rows[K7_NMS_CHUNK];
    int kcount;
    int pad[3];
    unsigned long long mask[K7_NMS_WORDS * K7_NMS_PITCH];
};

// Stable LSD radix sort (8 bits per pass) of n 64-bit keys by one CTA; returns the buffer holding the result.
__device__ uint64_t* k7_block_radix_sort(uint64_t* a, uint64_t* b, int n, K7SortSmem& S) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = pass * 8;
        if (tid < 256) S.hist[tid] = 0;
        if (tid == 0) S.skip = 0;
        __syncthreads();
        for (int i = tid; i < n; i += K7_NMS_THREADS) atomicAdd(&S.hist[(unsigned)(a[i] >> shift) & 255u], 1u);
        __syncthreads();
        if (tid < 256 && S.hist[tid] == (unsigned)n) S.skip = 1;   // every key has the same digit
        __syncthreads();
        if (S.skip) continue;
        if (warp < 8) {   // exclusive scan of the 256 bins: 8 warps x 32 bins
            const unsigned v = S.hist[tid];
            unsigned inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            S.base[tid] = inc - v;
            if (lane == 31) S.wtot[warp] = inc;
        }
        __syncthreads();
        if (tid < 256) {
            unsigned off = 0;
            for (int w = 0; w < (tid >> 5); ++w) off += S.wtot[w];
            S.base[tid] += off;
        }
        __syncthreads();
        for (int c0 = 0; c0 < n; c0 += K7_NMS_THREADS) {
            for (int k = tid; k < 32 * 256; k += K7_NMS_THREADS) (&S.warp_cnt[0][0])[k] = 0;
            __syncthreads();
            const int i = c0 + tid;
            const bool valid = i < n;
            const uint64_t key = valid ? a[i] : 0ull;
            const unsigned d = valid ? ((unsigned)(key >> shift) & 255u) : 256u;
            const unsigned peers = __match_any_sync(0xffffffffu, d);
            const int rank = __popc(peers & ((1u << lane) - 1u));
            if (valid && rank == 0) S.warp_cnt[warp][d] = __popc(peers);
            __syncthreads();
            if (tid < 256) {
                unsigned run = S.base[tid];
#pragma unroll 4
                for (int w = 0; w < 32; ++w) {
                    const unsigned c = S.warp_cnt[w][tid];
                    S.warp_cnt[w][tid] = run;
                    run += c;
                }
                S.base[tid] = run;
            }
            __syncthreads();
            if (valid) b[S.warp_cnt[warp][d] + rank] = key;
            __syncthreads();
        }
        uint64_t* t = a;
        a = b;
        b = t;
    }
    __syncthreads();
    return a;
}

// Bitonic sort of n <= K7_BITONIC_MAX keys in shared memory (keys are unique: the low word is the
// anchor id), result written to `out`.  ~log2(N)^2/2 barrier-separated steps; for the usual few
// hundred to few thousand survivors per image this replaces six to eight radix passes.
#define K7_BITONIC_MAX 4096
// sorts sk[0..N) ascending; N = power of two >= 64, entries beyond the n real keys hold ~0.
//
// Bitonic network (pair (i, i|j), ascending iff (i & k) == 0), run in as few barrier-separated
// passes as possible (the plain one-stage-per-barrier loop cost 91 barriers at N = 8192 and was
// half of k7_sort_kernel):
//   * a warp owns a block of 32*E consecutive keys, E per lane (key m of lane l = block[32m + l],
//     conflict-free): strides >= 32 inside the block are compare-exchanges between a lane's own
//     registers, strides < 32 are 64-bit shuffles with lane ^ j.  All phases k <= 32E (the complete
//     sort of every block) therefore run without a single barrier, and so do the last
//     log2(32E) stages of every later phase;
//   * the remaining strides (>= 32E) are taken up to three at a time: a thread loads the 8 keys
//     base + m*jl, runs the three stages in registers and writes them back.
// N = 8192: 13 barriers.
__device__ __forceinline__ void k7_cx(unsigned long long& lo, unsigned long long& hi, bool up) {
    const unsigned long long a = lo, b = hi;
    const bool sw = (a > b) == up;
    lo = sw ? b : a;
    hi = sw ? a : b;
}
// stages jstart, jstart/2, ..., 1 of phase k on one block; i0 = index in sk of this lane's key 0
template <int E>
__device__ __forceinline__ void k7_bitonic_block_stages(unsigned long long (&e)[E], int i0, int k, int jstart, int lane) {
#pragma unroll
    for (int dm = E / 2; dm >= 1; dm >>= 1) {
        if (dm * 32 > jstart) continue;
#pragma unroll
        for (int m = 0; m < E; ++m)
            if ((m & dm) == 0) k7_cx(e[m], e[m | dm], ((i0 + 32 * m) & k) == 0);
    }
#pragma unroll
    for (int j = 16; j >= 1; j >>= 1) {
        if (j > jstart) continue;
        const bool lower = (lane & j) == 0;
#pragma unroll
        for (int m = 0; m < E; ++m) {
            const unsigned long long o = __shfl_xor_sync(0xffffffffu, e[m], j);
            const bool up = ((i0 + 32 * m) & k) == 0;
            const bool take_min = lower == up;
            e[m] = ((o < e[m]) == take_min) ? o : e[m];
        }
    }
}
// stages j0, j0/2, ..., j0 >> (R-1) of phase k through shared memory, all strides >= 64
template <int R>
__device__ __forceinline__ void k7_bitonic_strided_pass(unsigned long long* sk, int N, int k, int j0) {
    const int jl = j0 >> (R - 1);
    for (int t = threadIdx.x; t < (N >> R); t += K7_NMS_THREADS) {
        const int low = t & (jl - 1);
        const int base = ((t - low) << R) | low;
        unsigned long long e[1 << R];
#pragma unroll
        for (int m = 0; m < (1 << R); ++m) e[m] = sk[base + m * jl];
        const bool up = (base & k) == 0;
#pragma unroll
        for (int dm = (1 << R) >> 1; dm >= 1; dm >>= 1) {
#pragma unroll
            for (int m = 0; m < (1 << R); ++m)
                if ((m & dm) == 0) k7_cx(e[m], e[m | dm], up);
        }
#pragma unroll
        for (int m = 0; m < (1 << R); ++m) sk[base + m * jl] = e[m];
    }
    __syncthreads();
}
template <int E>
__device__ void k7_bitonic_sort_smem_e(unsigned long long* sk, int N) {
    constexpr int BLK = 32 * E;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nblk = N / BLK;
    // every block sorted (alternating directions): phases k = 2 .. BLK
    for (int blk = warp; blk < nblk; blk += K7_NMS_THREADS / 32) {
        const int i0 = blk * BLK + lane;
        unsigned long long e[E];
#pragma unroll
        for (int m = 0; m < E; ++m) e[m] = sk[i0 + 32 * m];
        for (int k = 2; k <= BLK; k <<= 1) k7_bitonic_block_stages<E>(e, i0, k, k >> 1, lane);
#pragma unroll
        for (int m = 0; m < E; ++m) sk[i0 + 32 * m] = e[m];
    }
    __syncthreads();
    for (int k = 2 * BLK; k <= N; k <<= 1) {
        int j = k >> 1;
        while (j >= BLK) {   // strides through shared memory, three stages per pass
            if (j >= 4 * BLK) {
                k7_bitonic_strided_pass<3>(sk, N, k, j);
                j >>= 3;
            } else if (j >= 2 * BLK) {
                k7_bitonic_strided_pass<2>(sk, N, k, j);
                j >>= 2;
            } else {
                k7_bitonic_strided_pass<1>(sk, N, k, j);
                j >>= 1;
            }
        }
        for (int blk = warp; blk < nblk; blk += K7_NMS_THREADS / 32) {
            const int i0 = blk * BLK + lane;
            unsigned long long e[E];
#pragma unroll
            for (int m = 0; m < E; ++m) e[m] = sk[i0 + 32 * m];
            k7_bitonic_block_stages<E>(e, i0, k, BLK >> 1, lane);
#pragma unroll
            for (int m = 0; m < E; ++m) sk[i0 + 32 * m] = e[m];
        }
        __syncthreads();
    }
}
__device__ void k7_bitonic_sort_smem(unsigned long long* sk, int N) {
    // small sorts: 64-key blocks keep every warp busy; large ones: 256-key blocks, fewest barriers
    if (N >= 4096) k7_bitonic_sort_smem_e<8>(sk, N);
    else k7_bitonic_sort_smem_e<2>(sk, N);
}
__device__ __forceinline__ int k7_bitonic_size(int n) {
    int N = 64;
    while (N < n) N <<= 1;
    return N;
}
__device__ void k7_block_bitonic_sort(const uint64_t* __restrict__ in, uint64_t* __restrict__ out, int n,
                                      unsigned long long* sk) {
    const int tid = threadIdx.x;
    const int N = k7_bitonic_size(n);
    for (int i = tid; i < N; i += K7_NMS_THREADS) sk[i] = (i < n) ? in[i] : ~0ull;
    __syncthreads();
    k7_bitonic_sort_smem(sk, N);
    for (int i = tid; i < n; i += K7_NMS_THREADS) out[i] = sk[i];
    __syncthreads();
}

__device__ __forceinline__ float k7_key_score(uint64_t key) {
    uint32_t u = ~(uint32_t)(key >> 32);
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    return __uint_as_float(u);
}

// Greedy NMS over keys sorted in visiting order; box of key k is boxes[k & 0xffffffff].
// labels != nullptr (class of key k = labels[k & 0xffffffff]): only boxes of the same class suppress each other.
// Writes kept keys/boxes in keep order, returns the number kept.
__device__ int k7_block_nms(const uint64_t* __restrict__ keys, int n, const float4* __restrict__ boxes,
                            const int32_t* __restrict__ labels, float thr, int max_keep, uint64_t* __restrict__ kept_key,
                            float4* __restrict__ kept_box, K7NmsSmem& S) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int kept0 = 0;
    for (int c0 = 0; c0 < n && kept0 < max_keep; c0 += K7_NMS_CHUNK) {
        const int i = c0 + tid;
        const int cn = min(K7_NMS_CHUNK, n - c0);
        const bool valid = tid < cn;
        uint64_t key = 0;
        float4 bx = make_float4(0.f, 0.f, 0.f, 0.f);
        int mlabel = 0;
        if (valid) {
            key = keys[i];
            bx = __ldg(boxes + (uint32_t)(key & 0xffffffffull));
            if (labels) mlabel = __ldg(labels + (uint32_t)(key & 0xffffffffull));
        }
        const K7Box me = {bx.x, bx.y, bx.z, bx.w};
        const float area = k7_area(bx.x, bx.y, bx.z, bx.w);
        bool alive = valid;
        // (a) suppression by boxes kept in earlier chunks
        for (int k0 = 0; k0 < kept0; k0 += 256) {
            const int kn = min(256, kept0 - k0);
            __syncthreads();
            if (tid < kn) {
                const float4 kb = kept_box[k0 + tid];
                S.kbox[tid] = kb;
                S.karea[tid] = k7_area(kb.x, kb.y, kb.z, kb.w);
                S.klabel[tid] = labels ? __ldg(labels + (uint32_t)(kept_key[k0 + tid] & 0xffffffffull)) : 0;
            }
            __syncthreads();
            if (alive) {
                for (int k = 0; k < kn; ++k) {
                    const float4 kb = S.kbox[k];
                    const K7Box other = {kb.x, kb.y, kb.z, kb.w};
                    if (S.klabel[k] == mlabel && fminf(other.x2, me.x2) > fmaxf(other.x1, me.x1) &&
                        fminf(other.y2, me.y2) > fmaxf(other.y1, me.y1) && k7_nms_suppress(other, S.karea[k], me, area, thr)) {
                        alive = false;
                        break;
                    }
                }
            }
        }
        // (b) bitmask matrix of the chunk
        S.box[tid] = bx;
        S.area[tid] = area;
        S.label[tid] = mlabel;
        const unsigned ab = __ballot_sync(0xffffffffu, alive);
        if (lane == 0) S.alive_bits[warp] = ab;
        if (tid == 0) S.kcount = 0;
        __syncthreads();
        {
            // work items = (64-row block rb, mask word w >= rb, row r of the block): 64 consecutive
            // threads share one (rb, w) pair, i.e. read the same 64 column boxes (broadcast), and
            // every thread gets the same number of items whatever its row (the triangle is balanced)
            const int nw = (cn + 63) >> 6;
            const int items = (nw * (nw + 1) / 2) << 6;
            for (int q = tid; q < items; q += K7_NMS_THREADS) {
                int p = q >> 6, rb = 0;
                while (p >= nw - rb) {
                    p -= nw - rb;
                    ++rb;
                }
                const int w = rb + p;
                const int row = (rb << 6) | (q & 63);
                if (row >= cn || !((S.alive_bits[row >> 5] >> (row & 31)) & 1u)) continue;
                const float4 mb = S.box[row];
                const K7Box me2 = {mb.x, mb.y, mb.z, mb.w};
                const float marea = S.area[row];
                const int rlabel = S.label[row];
                unsigned long long word = 0ull;
                const int jbeg = max(w * 64, row + 1), jend = min(w * 64 + 64, cn);
                for (int j = jbeg; j < jend; ++j) {
                    const float4 ob = S.box[j];
                    const K7Box other = {ob.x, ob.y, ob.z, ob.w};
                    // an empty intersection gives overlap 0 (or NaN), never > thr >= 0: skip the division
                    if (S.label[j] == rlabel && fminf(me2.x2, other.x2) > fmaxf(me2.x1, other.x1) &&
                        fminf(me2.y2, other.y2) > fmaxf(me2.y1, other.y1) && k7_nms_suppress(me2, marea, other, S.area[j], thr))
                        word |= 1ull << (j & 63);
                }
                S.mask[w * K7_NMS_PITCH + row] = word;
            }
        }
        __syncthreads();
        // (c) greedy scan by warp 0, one 64-row block at a time.  Lane L holds the removed-word L.
        //     Inside a block the decision chain only needs the block's DIAGONAL words (64 x u64,
        //     two per lane, fetched before the chain starts and passed around with shuffles); the
        //     kept rows' other words are OR-ed into the later lanes afterwards with independent loads.
        if (warp == 0) {
            const int nw = (cn + 63) >> 6;
            unsigned long long remv = ~0ull;
            if (lane < K7_NMS_WORDS)
                remv = ~((unsigned long long)S.alive_bits[2 * lane] | ((unsigned long long)S.alive_bits[2 * lane + 1] << 32));
            int kc = 0;
            const int budget = max_keep - kept0;
            for (int w = 0; w < nw && kc < budget; ++w) {
                unsigned long long cur = __shfl_sync(0xffffffffu, remv, w);
                if (cur == ~0ull) continue;
                const int r0 = w << 6;
                const unsigned long long dlo = S.mask[w * K7_NMS_PITCH + r0 + lane];
                const unsigned long long dhi = S.mask[w * K7_NMS_PITCH + r0 + 32 + lane];
                unsigned long long keptm = 0ull;
#pragma unroll 8
                for (int r = 0; r < 64; ++r) {
                    const unsigned long long d = __shfl_sync(0xffffffffu, r < 32 ? dlo : dhi, r & 31);
                    if (!((cur >> r) & 1ull)) {   // warp-uniform
                        keptm |= 1ull << r;
                        cur |= d;
                    }
                }
                int nk = __popcll(keptm);
                if (kc + nk > budget) {   // keep only the first (budget - kc) rows of the block
                    nk = budget - kc;
                    unsigned long long t = keptm, first = 0ull;
                    for (int i = 0; i < nk; ++i) {
                        first |= t & (~t + 1ull);
                        t &= t - 1ull;
                    }
                    keptm = first;
                }
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int r = lane + 32 * h;
                    if ((keptm >> r) & 1ull) S.kept_rows[kc + __popcll(keptm & ((1ull << r) - 1ull))] = r0 + r;
                }
                if (lane > w && lane < nw) {
                    unsigned long long k = keptm;
                    while (k) {
                        const int r = __ffsll((long long)k) - 1;
                        k &= k - 1ull;
                        remv |= S.mask[lane * K7_NMS_PITCH + r0 + r];
                    }
                }
                kc += nk;
            }
            if (lane == 0) S.kcount = kc;
        }
        __syncthreads();
        const int kc = S.kcount;
        for (int k = tid; k < kc; k += K7_NMS_THREADS) {
            const int row = S.kept_rows[k];
            kept_key[kept0 + k] = keys[c0 + row];
            kept_box[kept0 + k] = S.box[row];
        }
        kept0 += kc;
        __syncthreads();
    }
    return kept0;
}

union K7SortNmsSmem {
    K7SortSmem sort;
    K7NmsSmem nms;
    unsigned long long bitonic[K7_BITONIC_MAX];
};

__device__ __forceinline__ uint64_t* k7_block_sort(uint64_t* a, uint64_t* b, int n, K7SortNmsSmem& S) {
    if (n <= K7_BITONIC_MAX) {
        k7_block_bitonic_sort(a, b, n, S.bitonic);
        return b;
    }
    return k7_block_radix_sort(a, b, n, S.sort);
}

// One image with more survivors than the multi-SM path is sized for (n > cap): radix / bitonic sort of
// the keys k7_sort_kernel left in keys0, then NMS in chunks of 1024 boxes, all by one CTA.  Called by
// k7_resolve_kernel for those images (no launch of its own: it is almost never needed).
__device__ void k7_fallback_image(int b, int A, float iou_thr, int pre_topk, int max_det,
                                  const int32_t* __restrict__ surv_count, int32_t* __restrict__ kept_count, uint64_t* keys0,
                                  uint64_t* keys1, const float4* __restrict__ box_ws, uint64_t* __restrict__ kept_key,
                                  float4* __restrict__ kept_box, int32_t* __restrict__ counts_out,
                                  const int32_t* __restrict__ label_ws, K7SortNmsSmem& S) {
    const int n_all = surv_count[b];
    const size_t ob = (size_t)b * A;
    uint64_t* sorted = k7_block_sort(keys0 + ob, keys1 + ob, n_all, S);
    int n = n_all;
    if (pre_topk > 0 && n > pre_topk) n = pre_topk;
    const int max_keep = max_det > 0 ? max_det : 0x7fffffff;
    const int kept = k7_block_nms(sorted, n, box_ws + ob, label_ws ? label_ws + ob : nullptr, iou_thr, max_keep,
                                  kept_key + ob, kept_box + ob, S.nms);
    if (threadIdx.x == 0) {
        kept_count[b] = kept;
        if (counts_out) {
            counts_out[2 * b + 0] = n_all;
            counts_out[2 * b + 1] = kept;
        }
    }
}

// ------------------------------------------------------------------ multi-SM NMS for n <= cap survivors
// Round-1 profiles: with one CTA per image the IoU bitmask (n^2/2 pair tests) and the greedy scan
// ran on 64 of 148 SMs and took longer than streaming the whole classification tensor; and at
// EfficientDet-D7 (7 500 survivors per image) even a bitmask built on every SM costs 28 M pair
// tests per image.  The NMS is therefore three kernels, and the bitmask is SPATIALLY indexed:
//   k7_sort_kernel   per image: survivor ranks (bitmap prefix), key sort by score (= the visiting
//                    order "rank"), then a second sort of the ranked boxes by (area class, x1) (=
//                    the "spatial index" p) with a summary (min x1, max x2, min/max area) of every
//                    block of 64 consecutive boxes;
//   k7_mask_kernel   grid (G, B): for every pair of 64-blocks that the summaries cannot rule out
//                    (x ranges overlap and the areas are within 1/thr of each other - IoU <= the
//                    area ratio), the 64 x 64 pair tests; row p gets the word of bits q "box q
//                    overlaps p above thr AND ranks before p".  Only non-zero words are written, each
//                    flagged in a per-row summary, so the n x n/64 matrix is never initialised or read
//                    as a whole;
//   k7_resolve_kernel  one CTA per image: greedy NMS as a parallel fixed point over those words.
// Which boxes share a 64-bit word is irrelevant to the result: every bit still says "a higher
// scored box overlaps me", so the kept set is exactly torchvision.ops.nms's.  Images with more
// survivors than the workspace's cap are served by k7_fallback_image (chunked, one CTA) inside k7_resolve_kernel.
#define K7_SORT_MAX 16384                       // largest cap: 128 KB of keys in shared memory
#define K7_NZW_MAX (K7_SORT_MAX / 64 / 64)      // u64 words of the per-row summary
#define K7_CELLS_MAX 4096                       // grid cells of the spatial order, all area binades together

// upper bound on survivors per image the multi-SM path is sized for (multiple of 64)
static int k7_nms_cap(int B, int n_max) {
    long long cap = ((long long)n_max + 63) / 64 * 64;
    if (cap > K7_SORT_MAX) cap = K7_SORT_MAX;
    // the bitmask is cap^2/8 bytes per image (sparsely touched): keep the batch under 4 GB
    const double budget = 4.0 * 1024 * 1024 * 1024;
    while (cap > 2048 && (double)cap * (double)cap / 8.0 * (double)B > budget) cap -= 64;
    return (int)cap;
}

// Per image: the survivor list is born here.  The score kernel only marked survivors in a bitmap
// (and left score / label / box by anchor id); this kernel takes the exclusive popcount prefix of
// the bitmap (= the index torchvision.ops.nms would return for each survivor), turns every set bit
// into a 64-bit key (score descending | anchor ascending) and sorts.
//   bits == nullptr: stand-alone NMS, keys0 already holds surv_count[b] keys.
// Outputs: surv_count[b]; keys1 = keys in rank order; sbox / srank / blk / nz in spatial order;
// nsel[b] = survivors the mask/resolve kernels must process, or -1 when the image has more than
// `cap` of them (then its unsorted keys are left in keys0 for the fallback).
__global__ void __launch_bounds__(K7_NMS_THREADS, 1)
k7_sort_kernel(int A, int W32, int pre_topk, int cap, int32_t* __restrict__ surv_count, const uint32_t* __restrict__ bits,
               int32_t* __restrict__ wprefix, const float* __restrict__ score_ws, uint64_t* __restrict__ keys0,
               uint64_t* __restrict__ keys1, const float4* __restrict__ box_ws, float4* __restrict__ sbox,
               int32_t* __restrict__ srank, float4* __restrict__ blk, unsigned long long* __restrict__ nz,
               int32_t* __restrict__ nsel, int32_t* __restrict__ work, const int32_t* __restrict__ label_ws,
               int32_t* __restrict__ slabel, float4* __restrict__ rbox) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned long long* sk = reinterpret_cast<unsigned long long*>(smem_raw);
    __shared__ unsigned int s_wtot[32];
    __shared__ float4 s_wsum[32];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t ob = (size_t)b * A;
    k7_pdl_trigger();   // the mask kernel's CTAs may take their places
    k7_pdl_wait();      // the score kernel's records
    int n_all;
    if (bits) {
        // Survivor ranks and keys.  A thread owns a contiguous chunk of the image's bitmap words:
        // one block scan gives every chunk its first rank.  The first walk over the bits also
        // prefetches what the gathers below will want (score and box by anchor id; the score
        // kernel wrote them long ago, most lines have left L2).
        const int chunk = (W32 + K7_NMS_THREADS - 1) / K7_NMS_THREADS;
        const int w_lo = min(tid * chunk, W32), w_hi = min(w_lo + chunk, W32);
        const uint32_t* brow = bits + (size_t)b * W32;
        int cnt = 0;
        for (int w = w_lo; w < w_hi; w += 4) {   // four words in flight (a thread's chunk is not coalesced)
            uint32_t wd[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) wd[u] = (w + u < w_hi) ? brow[w + u] : 0u;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                uint32_t word = wd[u];
                cnt += __popc(word);
                while (word) {
                    const int a = ((w + u) << 5) + __ffs(word) - 1;
                    word &= word - 1;
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(score_ws + ob + a));
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(box_ws + ob + a));
                }
            }
        }
        int inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_wtot[warp] = (unsigned)inc;
        __syncthreads();
        int slot = inc - cnt;
        n_all = 0;
        for (int k = 0; k < K7_NMS_THREADS / 32; ++k) {
            const int t = (int)s_wtot[k];
            if (k < warp) slot += t;
            n_all += t;
        }
        if (tid == 0) surv_count[b] = n_all;
        const bool small_ = n_all <= cap;
        // the anchor id of every set bit at its rank: shared memory (small image) or keys0 (fallback image)
        for (int w = w_lo; w < w_hi; w += 4) {
            uint32_t wd[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) wd[u] = (w + u < w_hi) ? brow[w + u] : 0u;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (w + u >= w_hi) break;
                uint32_t word = wd[u];
                wprefix[(size_t)b * W32 + w + u] = slot;
                while (word) {
                    const uint64_t a = (uint64_t)(((w + u) << 5) + __ffs(word) - 1);
                    word &= word - 1;
                    if (small_) sk[slot] = a; else keys0[ob + slot] = a;
                    ++slot;
                }
            }
        }
        __syncthreads();   // also: s_wtot is reused below
        // keys, every thread the same number of them with four score gathers in flight (walking the
        // bits and gathering in one loop left the whole CTA waiting for the thread with most bits)
        for (int i0 = 0; i0 < n_all; i0 += 4 * K7_NMS_THREADS) {
            uint32_t a[4];
            float sc[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * K7_NMS_THREADS + tid;
                if (i < n_all) {
                    a[u] = (uint32_t)(small_ ? sk[i] : keys0[ob + i]);
                    sc[u] = score_ws[ob + a[u]];
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int i = i0 + u * K7_NMS_THREADS + tid;
                if (i < n_all) {
                    const uint64_t key = ((uint64_t)k7_score_key_desc(sc[u]) << 32) | a[u];
                    if (small_) sk[i] = key; else keys0[ob + i] = key;
                }
            }
        }
    } else {
        n_all = surv_count[b];
    }
    const bool small = n_all <= cap;
    int N = k7_bitonic_size(n_all);
    if (bits) {
        if (small)
            for (int i = n_all + tid; i < N; i += K7_NMS_THREADS) sk[i] = ~0ull;
    } else if (small) {
        for (int i = tid; i < N; i += K7_NMS_THREADS) sk[i] = (i < n_all) ? keys0[ob + i] : ~0ull;
    }
    if (!small) {
        if (tid == 0) nsel[b] = -1;
        return;
    }
    __syncthreads();
    k7_bitonic_sort_smem(sk, N);   // ends with a barrier
    int n = n_all;
    if (pre_topk > 0 && n > pre_topk) n = pre_topk;
    if (tid == 0) nsel[b] = n;

    // ---- spatial order: boxes of one area binade, binned on a grid whose cells hold ~16 of them,
    // cells visited row by row in alternating direction (consecutive cells always touch).  Any
    // order gives the same NMS result; this one makes 64 consecutive boxes neighbours in the image
    // (about four cells in a row), so most pairs of 64-blocks are ruled out by their summaries in
    // k7_mask_kernel.
    // The boxes are gathered by anchor id ONCE (four independent loads in flight per thread) and
    // left in rank order in `rbox`; what the binning needs of a box - x1, y1 (28 bits each are
    // plenty for a cell index) and the binade of its area - replaces its key in sk[].
    __shared__ int s_cls[256], s_base[257], s_g[256];
    __shared__ int s_cell[K7_CELLS_MAX + 1];
    __shared__ float s_bnd[4];
    if (tid < 256) s_cls[tid] = 0;
    for (int c = tid; c <= K7_CELLS_MAX; c += K7_NMS_THREADS) s_cell[c] = 0;
    __syncthreads();
    float bx0 = 3.0e38f, bx1 = -3.0e38f, by0 = 3.0e38f, by1 = -3.0e38f;
    for (int i0 = 0; i0 < n; i0 += 4 * K7_NMS_THREADS) {
        uint64_t key[4];
        float4 bx[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * K7_NMS_THREADS + tid;
            if (i < n) {
                key[u] = sk[i];
                bx[u] = __ldg(box_ws + ob + (uint32_t)(key[u] & 0xffffffffull));
            }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * K7_NMS_THREADS + tid;
            if (i < n) {
                keys1[ob + i] = key[u];
                rbox[ob + i] = bx[u];
                const float area = k7_area(bx[u].x, bx[u].y, bx[u].z, bx[u].w);
                const unsigned c8 = (__float_as_uint(fmaxf(area, 0.0f)) >> 23) & 0xffu;   // binade of the area
                atomicAdd(&s_cls[c8], 1);
                bx0 = fminf(bx0, bx[u].x); bx1 = fmaxf(bx1, bx[u].x);
                by0 = fminf(by0, bx[u].y); by1 = fmaxf(by1, bx[u].y);
                sk[i] = ((unsigned long long)(__float_as_uint(bx[u].x) & 0xfffffff0u) << 32) |
                        ((unsigned long long)(__float_as_uint(bx[u].y) & 0xfffffff0u) << 4) | c8;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        bx0 = fminf(bx0, __shfl_xor_sync(0xffffffffu, bx0, o));
        bx1 = fmaxf(bx1, __shfl_xor_sync(0xffffffffu, bx1, o));
        by0 = fminf(by0, __shfl_xor_sync(0xffffffffu, by0, o));
        by1 = fmaxf(by1, __shfl_xor_sync(0xffffffffu, by1, o));
    }
    if (lane == 0) s_wsum[warp] = make_float4(bx0, bx1, by0, by1);
    __syncthreads();
    if (warp == 0) {
        float4 v = s_wsum[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            v.x = fminf(v.x, __shfl_xor_sync(0xffffffffu, v.x, o));
            v.y = fmaxf(v.y, __shfl_xor_sync(0xffffffffu, v.y, o));
            v.z = fminf(v.z, __shfl_xor_sync(0xffffffffu, v.z, o));
            v.w = fmaxf(v.w, __shfl_xor_sync(0xffffffffu, v.w, o));
        }
        if (lane == 0) {
            s_bnd[0] = v.x;
            s_bnd[1] = (v.y > v.x) ? 1.0f / (v.y - v.x) : 0.0f;
            s_bnd[2] = v.z;
            s_bnd[3] = (v.w > v.z) ? 1.0f / (v.w - v.z) : 0.0f;
        }
    }
    // grid side per binade: g*g cells of ~16 boxes; sum of g*g <= n/16 + 2*sqrt(16n) + 256 <= K7_CELLS_MAX
    if (tid < 256) {
        const int c = s_cls[tid];
        int g = 0;
        if (c > 0) {
            g = (int)ceilf(sqrtf((float)c * (1.0f / 16.0f)));
            g = max(1, min(g, 32));
        }
        s_g[tid] = g;
        int inc = g * g;
        const int v = inc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        s_base[tid] = inc - v;
        if (lane == 31) s_wtot[warp] = (unsigned)inc;
    }
    __syncthreads();
    if (tid < 256) {
        int off = 0;
        for (int w = 0; w < warp; ++w) off += (int)s_wtot[w];
        s_base[tid] += off;
    }
    __syncthreads();
    // cell of every box (kept in sk[i]) and the cell histogram
    for (int i = tid; i < n; i += K7_NMS_THREADS) {
        const unsigned long long v = sk[i];
        const float x = __uint_as_float((uint32_t)(v >> 32) & 0xfffffff0u);
        const float y = __uint_as_float((uint32_t)(v >> 4) & 0xfffffff0u);
        const int c8 = (int)(v & 0xffull);
        const int g = s_g[c8];
        int cx = (int)((x - s_bnd[0]) * s_bnd[1] * (float)g), cy = (int)((y - s_bnd[2]) * s_bnd[3] * (float)g);
        cx = max(0, min(cx, g - 1));   // also sends NaN to cell 0
        cy = max(0, min(cy, g - 1));
        if (cy & 1) cx = g - 1 - cx;   // odd rows run backwards
        const int cell = min(s_base[c8] + cy * g + cx, K7_CELLS_MAX - 1);
        sk[i] = (unsigned long long)cell;
        atomicAdd(&s_cell[cell], 1);
    }
    __syncthreads();
    // exclusive prefix over the cells (K7_CELLS_MAX = 4 * 1024 threads: four cells per thread)
    {
        static_assert(K7_CELLS_MAX == 4 * K7_NMS_THREADS, "four cells per thread");
        int c[4], inc = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            c[k] = s_cell[4 * tid + k];
            inc += c[k];
        }
        const int v = inc;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_wtot[warp] = (unsigned)inc;
        __syncthreads();
        int run = inc - v;
        for (int w = 0; w < warp; ++w) run += (int)s_wtot[w];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            s_cell[4 * tid + k] = run;
            run += c[k];
        }
    }
    __syncthreads();
    // scatter: the order inside a cell is whatever the atomics give (it does not affect the result)
    const size_t oc = (size_t)b * cap;
    const int nzw = (cap / 64 + 63) / 64;
    for (int i0 = 0; i0 < n; i0 += 4 * K7_NMS_THREADS) {
        float4 bx[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * K7_NMS_THREADS + tid;
            if (i < n) bx[u] = __ldcg(rbox + ob + i);   // written above by this thread
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * K7_NMS_THREADS + tid;
            if (i < n) {
                const int p = atomicAdd(&s_cell[(int)sk[i]], 1);
                sbox[oc + p] = bx[u];
                srank[oc + p] = i;
                if (slabel)   // class-aware NMS
                    slabel[oc + p] = __ldg(label_ws + ob + (uint32_t)(__ldcg(keys1 + ob + i) & 0xffffffffull));
            }
        }
    }
    {   // flags of the rows' non-zero words: only column-blocks < ceil(n/64) can be flagged
        const int nzu = ((n + 63) / 64 + 63) / 64;
        for (int i = tid; i < n * nzu; i += K7_NMS_THREADS) nz[(oc + i / nzu) * nzw + i % nzu] = 0ull;
    }
    if (tid == 0) work[b] = 0;
    __syncthreads();

    // ---- summaries of every 32 consecutive boxes: x range, y range, area range
    const int n32 = (n + 31) & ~31;
    for (int p0 = 0; p0 < n32; p0 += K7_NMS_THREADS) {
        const int p = p0 + tid;
        if (p0 + (warp << 5) >= n32) break;   // warp-uniform
        float x1 = 3.0e38f, x2 = -3.0e38f, y1 = 3.0e38f, y2 = -3.0e38f, amin = 3.0e38f, amax = -3.0e38f;
        if (p < n) {
            const float4 bx = __ldcg(sbox + oc + p);   // written by this CTA before the barrier
            const float area = k7_area(bx.x, bx.y, bx.z, bx.w);
            x1 = bx.x; y1 = bx.y; x2 = bx.z; y2 = bx.w; amin = area; amax = area;   // a NaN drops out of fminf/fmaxf
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            x1 = fminf(x1, __shfl_xor_sync(0xffffffffu, x1, o));
            x2 = fmaxf(x2, __shfl_xor_sync(0xffffffffu, x2, o));
            y1 = fminf(y1, __shfl_xor_sync(0xffffffffu, y1, o));
            y2 = fmaxf(y2, __shfl_xor_sync(0xffffffffu, y2, o));
            amin = fminf(amin, __shfl_xor_sync(0xffffffffu, amin, o));
            amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        }
        if (lane == 0) {
            float4* o = blk + 2 * ((size_t)b * (cap / 32) + (p >> 5));
            o[0] = make_float4(x1, x2, y1, y2);
            o[1] = make_float4(amin, amax, 0.0f, 0.0f);
        }
    }
}

// The exact decision of torchvision.ops.nms is fl(inter / (area_i + area_j - inter)) > thr.  The
// 64 x 64 pair loop only runs a cheap FILTER that no suppressed pair can fail:
//     w > 0, h > 0  and  inter > fl(c*area_i) + fl(c*area_j),   c = thr/(1+thr) * (1 - 1e-5).
// Proof: a suppressed pair has w, h > 0 (else inter is 0 or NaN and the quotient is not > thr >= 0),
// hence a positive union u, and fl(inter/u) > thr gives inter > thr*u*(1 - 2^-24), i.e.
// inter*(1 + thr') > thr'*(area_i + area_j)*(1 - 2^-23) with thr' = thr*(1 - 2^-23); the float32
// roundings of the two products and the sum are below 2^-22 relative, far inside the 1e-5 margin.
// Pairs that pass (about 1 in 50 overlapping ones) are re-tested exactly, division included.
__device__ __forceinline__ bool k7_nms_filter(const K7Box& a, float ca, const float4 b, float cb) {
    const float w = K7_FSUB(fminf(a.x2, b.z), fmaxf(a.x1, b.x));
    const float h = K7_FSUB(fminf(a.y2, b.w), fmaxf(a.y1, b.y));
    return (w > 0.0f) && (h > 0.0f) && (K7_FMUL(w, h) > K7_FADD(ca, cb));
}
__device__ __forceinline__ bool k7_nms_suppress_exact(const K7Box& a, float area_a, const float4 b, float area_b,
                                                      float thr) {
    const K7Box o = {b.x, b.y, b.z, b.w};
    return k7_nms_suppress(a, area_a, o, area_b, thr);
}

// Summary test of two sets of boxes: can they hold a pair with IoU > thr?  Exact in the "no"
// direction: (x, y) if min x1 of one set >= max x2 of the other (same for y), every pair has a
// non-positive intersection width -> overlap 0; (area) a pair with intersection > 0 has
// IoU <= min(area)/max(area), so sets whose areas differ by more than thr (1e-4 margin against
// float32 rounding) hold no pair above thr.  c_area == 0 (thr == 0) disables the area test.
struct K7Sum {
    float4 r;   // min x1, max x2, min y1, max y2
    float4 a;   // min area, max area
};
__device__ __forceinline__ K7Sum k7_sum_load(const float4* __restrict__ bs, int i32) {
    K7Sum s;
    s.r = __ldg(bs + 2 * i32);
    s.a = __ldg(bs + 2 * i32 + 1);
    return s;
}
__device__ __forceinline__ bool k7_sum_may_overlap(const K7Sum& p, const K7Sum& q, float c_area) {
    return (q.r.x < p.r.y) && (p.r.x < q.r.y) && (q.r.z < p.r.w) && (p.r.z < q.r.w) &&
           !(q.a.y < c_area * p.a.x) && !(p.a.y < c_area * q.a.x);
}

// A group of 64 threads (two warps) works on a row-block: thread t keeps box p = 64*rb + t in
// registers.  Candidate column-blocks cb <= rb are found with the 32-box summaries (64 at a time,
// one per thread, ballots); for every candidate the group stages the 64 column boxes in shared
// memory and each warp runs its 32 x 32 sub-blocks that the summaries do not rule out: 32 or 64
// filter tests per thread, exact re-test of the few pairs that pass.  Every unordered pair is
// tested ONCE: an overlap found by thread p sets bit q in row p when q ranks before p, otherwise
// bit p in row q (collected per column in shared memory and written by thread q afterwards).
#define K7_MASK_THREADS 256
#define K7_MASK_GROUPS (K7_MASK_THREADS / 64)
__global__ void __launch_bounds__(K7_MASK_THREADS)
k7_mask_kernel(float thr, int cap, const int32_t* __restrict__ nsel, const float4* __restrict__ sbox,
               const int32_t* __restrict__ srank, const float4* __restrict__ blk, unsigned long long* __restrict__ nz,
               unsigned long long* __restrict__ M, int32_t* __restrict__ work, const int32_t* __restrict__ slabel) {
    __shared__ float4 cbox[K7_MASK_GROUPS][64];
    __shared__ float carea[K7_MASK_GROUPS][64];
    __shared__ int crank[K7_MASK_GROUPS][64];
    __shared__ int clabel[K7_MASK_GROUPS][64];
    __shared__ unsigned long long tword[K7_MASK_GROUPS][64];
    __shared__ unsigned int cand[K7_MASK_GROUPS][2];
    __shared__ int s_item[K7_MASK_GROUPS];
    const int b = blockIdx.y;
    k7_pdl_trigger();
    k7_pdl_wait();      // k7_sort_kernel's spatial order
    const int n = nsel[b];
    if (n <= 0) return;
    const int nb = (n + 63) >> 6, n32 = (n + 31) >> 5;
    const int nbw = cap >> 6, nzw = (nbw + 63) >> 6;
    const int grp = threadIdx.x >> 6, t = threadIdx.x & 63, lane = threadIdx.x & 31, half = t >> 5;
    const float c_flt = thr / (1.0f + thr) * (1.0f - 1e-5f);   // filter constant, see k7_nms_filter
    const float c_area = thr * (1.0f - 1e-4f);
    const size_t oc = (size_t)b * cap;
    const float4* bx = sbox + oc;
    const float4* bs = blk + 2 * (size_t)b * (cap >> 5);
    const int bar_id = 1 + grp;
    // work item = (row-block, share of its candidate column-blocks), handed out by a per-image counter.
    // At least two items per group of the image's CTAs: with few images per launch (a sharded
    // batch) the row-blocks alone leave most SMs idle and the heavy ones become the tail.
    int csplit = nb >= 64 ? 2 : (nb >= 16 ? 4 : 8);
    if (nb >= 32) {
        const int want = (2 * (int)gridDim.x * K7_MASK_GROUPS + nb - 1) / nb;
        csplit = max(csplit, min(want, 16));
    }
    const int items = nb * csplit;
    for (;;) {
        if (t == 0) s_item[grp] = atomicAdd(work + b, 1);
        asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
        const int wi = s_item[grp];
        asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
        if (wi >= items) break;   // group-uniform
        // heavy row-blocks (many column-blocks below them) first
        const int rb = nb - 1 - wi / csplit, sub = wi % csplit;
        int turn = 0, mine_turn = sub;
        const int row = rb * 64 + t;
        const bool rlive = row < n;
        const float4 mb = rlive ? __ldg(bx + row) : make_float4(0.f, 0.f, 0.f, 0.f);
        const K7Box me = {mb.x, mb.y, mb.z, mb.w};
        const float marea = k7_area(mb.x, mb.y, mb.z, mb.w);
        const float mca = K7_FMUL(c_flt, marea);
        const int mrank = rlive ? __ldg(srank + oc + row) : -1;
        const int mlabel = (slabel && rlive) ? __ldg(slabel + oc + row) : 0;
        // summaries of the two halves of the row-block; rmine = the half this warp owns
        const K7Sum r0 = k7_sum_load(bs, 2 * rb);
        const bool has_r1 = 2 * rb + 1 < n32;
        const K7Sum r1 = has_r1 ? k7_sum_load(bs, 2 * rb + 1) : r0;
        const K7Sum rmine = half ? r1 : r0;
        for (int c0 = 0; c0 <= rb; c0 += 64) {
            // which of column-blocks c0..c0+63 (<= rb) can hold a pair above thr with this row-block?
            const int cbt = c0 + t;
            bool ok = false;
            if (cbt <= rb) {
                const K7Sum q0 = k7_sum_load(bs, 2 * cbt);
                ok = k7_sum_may_overlap(r0, q0, c_area) || (has_r1 && k7_sum_may_overlap(r1, q0, c_area));
                if (2 * cbt + 1 < n32) {
                    const K7Sum q1 = k7_sum_load(bs, 2 * cbt + 1);
                    ok = ok || k7_sum_may_overlap(r0, q1, c_area) || (has_r1 && k7_sum_may_overlap(r1, q1, c_area));
                }
            }
            const unsigned bal = __ballot_sync(0xffffffffu, ok);
            if (lane == 0) cand[grp][half] = bal;
            asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
            unsigned long long cm = (unsigned long long)cand[grp][0] | ((unsigned long long)cand[grp][1] << 32);
            while (cm) {   // group-uniform
                const int cb = c0 + __ffsll((long long)cm) - 1;
                cm &= cm - 1ull;
                if (turn++ != mine_turn) continue;   // every csplit-th candidate is this item's (no modulo in the loop)
                mine_turn += csplit;
                const int col = cb * 64 + t;
                const bool diag = cb == rb;
                asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");   // previous column-block fully consumed
                if (col < n) {
                    const float4 c = __ldg(bx + col);
                    cbox[grp][t] = c;
                    carea[grp][t] = K7_FMUL(c_flt, k7_area(c.x, c.y, c.z, c.w));
                    crank[grp][t] = __ldg(srank + oc + col);
                    clabel[grp][t] = slabel ? __ldg(slabel + oc + col) : 0;
                }
                tword[grp][t] = 0ull;
                asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
                const int jend = min(64, n - cb * 64);
                if (rlive && (!half || has_r1)) {
                    uint32_t flo = 0u, fhi = 0u;   // pairs that pass the filter
                    // warp-uniform: this warp's 32 rows against each half of the column-block
                    if (k7_sum_may_overlap(rmine, k7_sum_load(bs, 2 * cb), c_area)) {
                        const int je = min(32, jend);
                        if (je == 32) {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (k7_nms_filter(me, mca, cbox[grp][j], carea[grp][j])) flo |= 1u << j;
                        } else {
                            for (int j = 0; j < je; ++j)
                                if (k7_nms_filter(me, mca, cbox[grp][j], carea[grp][j])) flo |= 1u << j;
                        }
                    }
                    if (jend > 32 && k7_sum_may_overlap(rmine, k7_sum_load(bs, 2 * cb + 1), c_area)) {
                        if (jend == 64) {
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (k7_nms_filter(me, mca, cbox[grp][j + 32], carea[grp][j + 32])) fhi |= 1u << j;
                        } else {
                            for (int j = 32; j < jend; ++j)
                                if (k7_nms_filter(me, mca, cbox[grp][j], carea[grp][j])) fhi |= 1u << (j - 32);
                        }
                    }
                    unsigned long long f = (unsigned long long)flo | ((unsigned long long)fhi << 32);
                    unsigned long long word = 0ull;
                    while (f) {   // exact test; the box visited earlier suppresses the other one
                        const int j = __ffsll((long long)f) - 1;
                        f &= f - 1ull;
                        const float4 c = cbox[grp][j];
                        if (clabel[grp][j] != mlabel) continue;   // class-aware: other classes never suppress
                        if (!k7_nms_suppress_exact(me, marea, c, k7_area(c.x, c.y, c.z, c.w), thr)) continue;
                        if (crank[grp][j] < mrank)
                            word |= 1ull << j;
                        else if (!diag)   // in a diagonal block thread q meets this pair itself
                            atomicOr(reinterpret_cast<unsigned int*>(&tword[grp][j]) + (t >> 5), 1u << (t & 31));   // native 32-bit ATOMS
                    }
                    if (word) {
                        M[(oc + row) * nbw + cb] = word;
                        atomicOr(nz + (oc + row) * nzw + (cb >> 6), 1ull << (cb & 63));
                    }
                }
                if (!diag) {   // group-uniform: the transposed bits, one column box per thread
                    asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");
                    const unsigned long long tw = tword[grp][t];
                    if (tw) {
                        M[(oc + col) * nbw + rb] = tw;
                        atomicOr(nz + (oc + col) * nzw + (rb >> 6), 1ull << (rb & 63));
                    }
                }
            }
            asm volatile("bar.sync %0, 64;" ::"r"(bar_id) : "memory");   // cand[] is rewritten by the next round
        }
    }
}

// Greedy NMS as a fixed point, one CTA per image.  Row p holds the boxes q that overlap box p
// above the threshold and rank before it.  Box p is REMOVED as soon as one of them is known KEPT,
// and KEPT as soon as all of them are known removed; the undecided box of lowest rank can always be
// decided, so the iteration ends, and by induction over the rank its result is exactly the
// sequential greedy answer of torchvision.ops.nms.  The kept/removed sets only grow, so threads may
// read them while others update them; a round costs one barrier.  Dependency chains between
// overlapping boxes are short, so a handful of rounds replace the n-step sequential scan; rows
// without a flagged word (most) are kept in round 0 without touching the matrix.
#define K7_RESOLVE_THREADS 1024
#define K7_RESOLVE_PER (K7_SORT_MAX / K7_RESOLVE_THREADS)
// set bit p of a bit set kept as u64 words in shared memory.  A 64-bit atomicOr on shared memory
// is a compare-and-swap loop (ncu: 45 % of the old kernel's instructions, 64 threads fighting over
// one word); the 32-bit form is one native ATOMS.OR on the half that holds the bit (little endian).
__device__ __forceinline__ void k7_set_bit(unsigned long long* set, int p) {
    atomicOr(reinterpret_cast<unsigned int*>(set) + (p >> 5), 1u << (p & 31));
}
// The flagged words of an image (index + 64 bits, 10 bytes) are staged ONCE in shared memory:
// a round then costs shared-memory latency instead of two dependent trips to L2 per undecided
// row (ncu at EfficientDet-D7: 77 us per image, all of it long-scoreboard and barrier stalls).
// Images with more flagged words than this, or a row with more than 15 of them, iterate on
// the matrix in global memory as before.
#define K7_RESOLVE_LIST 12288
#define K7_RESOLVE_LIST_BYTES (K7_RESOLVE_LIST * (sizeof(unsigned long long) + sizeof(unsigned short)))
#define K7_RESOLVE_SMEM (K7_RESOLVE_LIST_BYTES > sizeof(K7SortNmsSmem) ? K7_RESOLVE_LIST_BYTES : sizeof(K7SortNmsSmem))
__global__ void __launch_bounds__(K7_RESOLVE_THREADS)
k7_resolve_kernel(int A, int max_det, int cap, const int32_t* __restrict__ nsel, const int32_t* __restrict__ surv_count,
                  uint64_t* sorted_keys, const float4* __restrict__ sbox,
                  const int32_t* __restrict__ srank, const unsigned long long* __restrict__ nz,
                  const unsigned long long* __restrict__ M, uint64_t* __restrict__ kept_key, float4* __restrict__ kept_box,
                  int32_t* __restrict__ kept_count, int32_t* __restrict__ counts_out, int64_t* __restrict__ keep_idx_out,
                  float iou_thr, int pre_topk, uint64_t* keys0, const float4* __restrict__ box_ws,
                  const int32_t* __restrict__ label_ws) {
    constexpr int WMAX = K7_SORT_MAX / 64;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned long long* s_l = reinterpret_cast<unsigned long long*>(smem_raw);
    unsigned short* s_w = reinterpret_cast<unsigned short*>(s_l + K7_RESOLVE_LIST);
    __shared__ unsigned long long s_kept[WMAX], s_rem[WMAX], s_krank[WMAX];
    __shared__ int s_prefix[WMAX + 1];
    __shared__ int s_wtot[32];
    __shared__ int s_undecided;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    k7_pdl_trigger();
    k7_pdl_wait();      // k7_mask_kernel's words
    const int n = nsel[b];
    if (n < 0) {   // more survivors than the multi-SM path is sized for
        k7_fallback_image(b, A, iou_thr, pre_topk, max_det, surv_count, kept_count, keys0, sorted_keys, box_ws, kept_key,
                          kept_box, counts_out, label_ws, *reinterpret_cast<K7SortNmsSmem*>(smem_raw));
        return;
    }
    const size_t ob = (size_t)b * A, oc = (size_t)b * cap;
    const int nbw = cap >> 6, nzw = (nbw + 63) >> 6;
    const int nw = (n + 63) >> 6;
    const int nzu = (nw + 63) >> 6;   // summary words that can hold a flag: column-blocks < nw
    for (int w = tid; w < WMAX; w += K7_RESOLVE_THREADS) s_kept[w] = s_rem[w] = s_krank[w] = 0ull;
    __syncthreads();
    uint32_t open_ = 0u;                 // bit k: row tid + k*1024 is still undecided
    unsigned long long cnts = 0ull;      // 4 bits per k: flagged words of that row (saturating at 15)
    int mine = 0;
    bool big = false;
#pragma unroll
    for (int k = 0; k < K7_RESOLVE_PER; ++k) {   // independent loads, no atomics in between
        if (k * K7_RESOLVE_THREADS >= n) break;   // block-uniform: no row of this image left
        const int p = tid + k * K7_RESOLVE_THREADS;
        if (p >= n) continue;
        int c = 0;
        for (int z = 0; z < nzu; ++z) c += __popcll(__ldcg(nz + (oc + p) * nzw + z));
        if (c > 0) {
            open_ |= 1u << k;
            big |= c > 15;
            cnts |= (unsigned long long)min(c, 15) << (4 * k);
            mine += c;
        }
    }
#pragma unroll
    for (int k = 0; k < K7_RESOLVE_PER; ++k) {
        if (k * K7_RESOLVE_THREADS >= n) break;   // block-uniform: no row of this image left
        // nothing ranked before it overlaps: kept.  A warp owns 32 consecutive rows = one 32-bit half-word.
        const int p = tid + k * K7_RESOLVE_THREADS;
        const unsigned kept0 = __ballot_sync(0xffffffffu, p < n && !((open_ >> k) & 1u));
        if (lane == 0 && kept0) reinterpret_cast<unsigned int*>(s_kept)[p >> 5] = kept0;
    }
    // exclusive scan of the per-thread entry counts
    int inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_wtot[warp] = inc;
    const int any_big = __syncthreads_or(big ? 1 : 0);
    int off = inc - mine, total = 0;
    for (int w = 0; w < K7_RESOLVE_THREADS / 32; ++w) {
        const int t = s_wtot[w];
        if (w < warp) off += t;
        total += t;
    }
    const bool staged = !any_big && total <= K7_RESOLVE_LIST;
    const uint32_t open0 = open_;
    if (staged) {
        int e = off;
        uint32_t o = open0;
        while (o) {
            const int k = __ffs(o) - 1;
            o &= o - 1u;
            const int p = tid + k * K7_RESOLVE_THREADS;
            const unsigned long long* row = M + (oc + p) * nbw;
            for (int z = 0; z < nzu; ++z) {
                unsigned long long m = __ldcg(nz + (oc + p) * nzw + z);
                while (m) {
                    const int w = (z << 6) + __ffsll((long long)m) - 1;
                    m &= m - 1ull;
                    s_w[e] = (unsigned short)w;
                    s_l[e] = __ldcg(row + w);
                    ++e;
                }
            }
        }
    }
    for (int round = 0; round < K7_SORT_MAX; ++round) {
        if (tid == 0) s_undecided = 0;
        __syncthreads();
        if (staged) {
            int e = off;
            uint32_t o = open0;
            while (o) {
                const int k = __ffs(o) - 1;
                o &= o - 1u;
                const int c = (int)((cnts >> (4 * k)) & 15ull);
                if ((open_ >> k) & 1u) {
                    bool any_kept = false, all_removed = true;
                    for (int i = 0; i < c; ++i) {
                        const int w = s_w[e + i];
                        const unsigned long long l = s_l[e + i];
                        const unsigned long long kp = *((volatile unsigned long long*)&s_kept[w]);
                        const unsigned long long rm = *((volatile unsigned long long*)&s_rem[w]);
                        any_kept |= (l & kp) != 0ull;
                        all_removed &= (l & ~rm) == 0ull;
                    }
                    const int p = tid + k * K7_RESOLVE_THREADS;
                    if (any_kept) {
                        k7_set_bit(s_rem, p);
                        open_ &= ~(1u << k);
                    } else if (all_removed) {
                        k7_set_bit(s_kept, p);
                        open_ &= ~(1u << k);
                    }
                }
                e += c;
            }
        } else {
            uint32_t o = open_;
            while (o) {
                const int k = __ffs(o) - 1;
                o &= o - 1u;
                const int p = tid + k * K7_RESOLVE_THREADS;
                const unsigned long long* row = M + (oc + p) * nbw;
                bool any_kept = false, all_removed = true;
                for (int z = 0; z < nzu; ++z) {
                    unsigned long long m = __ldcg(nz + (oc + p) * nzw + z);
                    while (m) {
                        const int w = (z << 6) + __ffsll((long long)m) - 1;
                        m &= m - 1ull;
                        const unsigned long long l = __ldcg(row + w);
                        const unsigned long long kp = *((volatile unsigned long long*)&s_kept[w]);
                        const unsigned long long rm = *((volatile unsigned long long*)&s_rem[w]);
                        any_kept |= (l & kp) != 0ull;
                        all_removed &= (l & ~rm) == 0ull;
                    }
                }
                if (any_kept) {
                    k7_set_bit(s_rem, p);
                    open_ &= ~(1u << k);
                } else if (all_removed) {
                    k7_set_bit(s_kept, p);
                    open_ &= ~(1u << k);
                }
            }
        }
        if (open_) s_undecided = 1;
        __syncthreads();
        if (s_undecided == 0) break;
        __syncthreads();
    }
    // kept set by rank, then compaction in rank (= visiting) order
#pragma unroll
    for (int k = 0; k < K7_RESOLVE_PER; ++k) {
        if (k * K7_RESOLVE_THREADS >= n) break;   // block-uniform: no row of this image left
        const int p = tid + k * K7_RESOLVE_THREADS;
        if (p >= n) continue;
        if ((s_kept[p >> 6] >> (p & 63)) & 1ull) {
            const int r = srank[oc + p];
            k7_set_bit(s_krank, r);
        }
    }
    __syncthreads();
    if (tid < 32) {   // exclusive prefix of the popcounts of s_krank[0..nw)
        int carry = 0;
        for (int w0 = 0; w0 < nw; w0 += 32) {
            const int w = w0 + tid;
            const int v = (w < nw) ? __popcll(s_krank[w]) : 0;
            int inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (tid >= o) inc += t;
            }
            if (w < nw) s_prefix[w] = carry + inc - v;
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (tid == 0) s_prefix[nw] = carry;
    }
    __syncthreads();
    const int budget = max_det > 0 ? max_det : 0x7fffffff;
#pragma unroll
    for (int k = 0; k < K7_RESOLVE_PER; ++k) {
        if (k * K7_RESOLVE_THREADS >= n) break;   // block-uniform: no row of this image left
        const int p = tid + k * K7_RESOLVE_THREADS;
        if (p >= n) continue;
        if (!((s_kept[p >> 6] >> (p & 63)) & 1ull)) continue;
        const int r = srank[oc + p];
        const int pos = s_prefix[r >> 6] + __popcll(s_krank[r >> 6] & ((1ull << (r & 63)) - 1ull));
        if (pos >= budget) continue;
        const uint64_t key = sorted_keys[ob + r];
        kept_key[ob + pos] = key;
        kept_box[ob + pos] = sbox[oc + p];
        if (keep_idx_out) keep_idx_out[pos] = (int64_t)(uint32_t)(key & 0xffffffffull);
    }
    if (tid == 0) {
        const int kc = min(s_prefix[nw], budget);
        kept_count[b] = kc;
        if (counts_out) {
            counts_out[2 * b + 0] = surv_count[b];
            counts_out[2 * b + 1] = kc;
        }
    }
}

// ------------------------------------------------------------------ gather
K7V1Ws workspace(void* base, int B, int A) {
    K7V1Ws w;
    unsigned char* p = static_cast<unsigned char*>(base);
    size_t o = 0;
    auto take = [&](size_t bytes) {
        unsigned char* r = p ? p + o : nullptr;
        o += k7_align_up(bytes, 256);
        return r;
    };
    w.cap = k7_nms_cap(B, A);
    const size_t cap = (size_t)w.cap, nbw = cap / 64, nzw = (nbw + 63) / 64;
    w.sbox = reinterpret_cast<float4*>(take(sizeof(float4) * (size_t)B * cap));
    w.srank = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * (size_t)B * cap));
    w.slabel = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * (size_t)B * cap));
    w.blk = reinterpret_cast<float4*>(take(sizeof(float4) * 4 * (size_t)B * nbw));
    w.work = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * (size_t)B));
    w.nz = reinterpret_cast<unsigned long long*>(take(8 * (size_t)B * cap * nzw));
    w.M = reinterpret_cast<unsigned long long*>(take(8 * (size_t)B * cap * nbw));
    w.bytes = o;
    return w;
}

int chain(int B, int A, float iou_thr, int pre_topk, int max_det, bool ca, int32_t* surv_count, uint64_t* keys0,
          uint64_t* keys1, const float4* box, const int32_t* label, int32_t* nsel, uint64_t* kept_key, float4* kept_box,
          int32_t* kept_count, int32_t* counts, const K7V1Ws& w, cudaStream_t st) {
    size_t smem_sort = 64 * 8;
    while (smem_sort < (size_t)w.cap * 8) smem_sort <<= 1;   // bitonic size of the largest image
    K7_CUDA(cudaFuncSetAttribute(k7_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_sort));
    k7_pdl_site() = 0;
    // bits == nullptr: the keys are already in keys0 (the score kernel wrote them with the survivor count)
    K7_CUDA(k7_launch_pdl(k7_sort_kernel, dim3(B), dim3(K7_NMS_THREADS), smem_sort, st, A, 0, pre_topk, w.cap, surv_count,
                          (const uint32_t*)nullptr, (int32_t*)nullptr, (const float*)nullptr, keys0, keys1, box, w.sbox, w.srank,
                          w.blk, w.nz, nsel, w.work, ca ? label : (const int32_t*)nullptr, ca ? w.slabel : (int32_t*)nullptr,
                          kept_box));
    int gx = (8 * k7_sm_count() + B - 1) / B;   // about eight CTAs per SM over the batch
    const int gx_max = (w.cap / 64 + K7_MASK_GROUPS - 1) / K7_MASK_GROUPS * 4;
    if (gx > gx_max) gx = gx_max;
    if (gx < 1) gx = 1;
    k7_pdl_site() = 1;
    K7_CUDA(k7_launch_pdl(k7_mask_kernel, dim3(gx, B), dim3(K7_MASK_THREADS), 0, st, iou_thr, w.cap, (const int32_t*)nsel,
                          (const float4*)w.sbox, (const int32_t*)w.srank, (const float4*)w.blk, w.nz, w.M, w.work,
                          ca ? (const int32_t*)w.slabel : (const int32_t*)nullptr));
    K7_CUDA(cudaFuncSetAttribute(k7_resolve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K7_RESOLVE_SMEM));
    k7_pdl_site() = 2;
    K7_CUDA(k7_launch_pdl(k7_resolve_kernel, dim3(B), dim3(K7_RESOLVE_THREADS), K7_RESOLVE_SMEM, st, A, max_det, w.cap, nsel,
                          surv_count, keys1, w.sbox, w.srank, w.nz, w.M, kept_key, kept_box, kept_count, counts,
                          (int64_t*)nullptr, iou_thr, pre_topk, keys0, box, ca ? label : (const int32_t*)nullptr));
    return K7_OK;
}

}  // namespace k7v1
