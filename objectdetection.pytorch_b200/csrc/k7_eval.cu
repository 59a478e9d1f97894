// Evaluator tails on device: what datasettool/coco_eval.py:25-57,
// ksevendata_eval.py:24-61 and csv_eval.py:93-121 do per image on the host, for a whole batch.
// Three small launches: count (one CTA per image) -> scan (one CTA) -> write (one CTA per image).
// Detections per image are few (tens to a few thousand): the point is not bandwidth but removing
// the per-image device->host copies and Python loops; the result is one packed float64 [N,7] array.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/k7det.h"
#include "k7_common.cuh"

#define K7_EVAL_THREADS 256

struct K7EvalDev {
    const float* scores;
    const int64_t* labels;
    const float* boxes;
    const int32_t* counts;
    const float* scale;
    const int64_t* image_ids;
    const int32_t* category_of_label;
    const uint8_t* valid_label;
    int cap, counts_stride, B, mode, max_det, nc;
    double thr;
    float thr_f;
};

// Exclusive rank of `flag` among the block's threads in thread order, and the block total.
// s_w: K7_EVAL_THREADS / 32 ints.  Two barriers.
__device__ __forceinline__ int k7_block_rank(const bool flag, int* s_w, int& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned bal = __ballot_sync(0xffffffffu, flag);
    __syncthreads();   // s_w may still be read from the previous use
    if (lane == 0) s_w[warp] = __popc(bal);
    __syncthreads();
    int base = 0;
    total = 0;
#pragma unroll
    for (int w = 0; w < K7_EVAL_THREADS / 32; ++w) {
        const int c = s_w[w];
        if (w < warp) base += c;
        total += c;
    }
    return base + __popc(bal & ((1u << lane) - 1u));
}

__device__ __forceinline__ bool k7_eval_label_ok(const K7EvalDev& E, const long long label) {
    if (label < 0 || label >= (long long)E.nc) return false;
    return E.mode == K7_EVAL_CSV || !E.valid_label || E.valid_label[label] != 0;
}

// COCO: the first row (in order) of an accepted class whose float(score) < thr ends the image
// (coco_eval.py:45-46; ksevendata_eval.py:46-51 tests the class first).  Returns that index (n if none).
__device__ int k7_eval_coco_stop(const K7EvalDev& E, const int b, const int n, int* s_min) {
    if (threadIdx.x == 0) *s_min = n;
    __syncthreads();
    int f = n;
    for (int i = threadIdx.x; i < n; i += K7_EVAL_THREADS) {
        const size_t o = (size_t)b * E.cap + i;
        if (k7_eval_label_ok(E, E.labels[o]) && (double)E.scores[o] < E.thr) {
            f = i;
            break;   // later rows of this thread are larger
        }
    }
    if (f < n) atomicMin(s_min, f);
    __syncthreads();
    return *s_min;
}

// kernel 1: rows emitted per image; CSV: also per (image, label).  Dynamic smem: nc ints (CSV).
__global__ void __launch_bounds__(K7_EVAL_THREADS)
k7_eval_count_kernel(const __grid_constant__ K7EvalDev E, int32_t* __restrict__ img_count, int32_t* __restrict__ class_count) {
    extern __shared__ int s_hist[];
    __shared__ int s_w[K7_EVAL_THREADS / 32];
    __shared__ int s_min;
    const int b = blockIdx.x;
    const int n = max(0, min(E.counts[(size_t)b * E.counts_stride], E.cap));
    int emitted = 0;
    if (E.mode == K7_EVAL_COCO) {
        const int f = k7_eval_coco_stop(E, b, n, &s_min);
        for (int i0 = 0; i0 < f; i0 += K7_EVAL_THREADS) {
            const int i = i0 + threadIdx.x;
            const bool keep = i < f && k7_eval_label_ok(E, E.labels[(size_t)b * E.cap + i]);
            int total;
            k7_block_rank(keep, s_w, total);
            emitted += total;
        }
    } else {
        for (int c = threadIdx.x; c < E.nc; c += K7_EVAL_THREADS) s_hist[c] = 0;
        __syncthreads();
        for (int i0 = 0; i0 < n && (E.max_det <= 0 || emitted < E.max_det); i0 += K7_EVAL_THREADS) {
            const int i = i0 + threadIdx.x;
            const size_t o = (size_t)b * E.cap + (i < n ? i : 0);
            const long long label = E.labels[o];
            const bool sel = i < n && E.scores[o] > E.thr_f && k7_eval_label_ok(E, label);
            int total;
            const int r = emitted + k7_block_rank(sel, s_w, total);
            if (sel && (E.max_det <= 0 || r < E.max_det)) atomicAdd(s_hist + (int)label, 1);
            emitted += total;
        }
        if (E.max_det > 0 && emitted > E.max_det) emitted = E.max_det;
        __syncthreads();
        for (int c = threadIdx.x; c < E.nc; c += K7_EVAL_THREADS) class_count[(size_t)b * E.nc + c] = s_hist[c];
    }
    if (threadIdx.x == 0) img_count[b] = emitted;
}

// kernel 2: exclusive scans (one CTA): img_count[B] -> row_offsets[B+1]; class_count[L] -> class_offsets[L+1].
__device__ void k7_block_scan_array(const int32_t* __restrict__ in, int32_t* __restrict__ out, const int L, int* s_part) {
    const int T = blockDim.x;
    const int per = (L + T - 1) / T;
    const int lo = min(L, (int)threadIdx.x * per), hi = min(L, lo + per);
    int sum = 0;
    for (int i = lo; i < hi; ++i) sum += in[i];
    __syncthreads();
    s_part[threadIdx.x] = sum;
    __syncthreads();
    for (int d = 1; d < T; d <<= 1) {   // Hillis-Steele inclusive scan of the partials
        const int v = (int)threadIdx.x >= d ? s_part[threadIdx.x - d] : 0;
        __syncthreads();
        s_part[threadIdx.x] += v;
        __syncthreads();
    }
    int run = s_part[threadIdx.x] - sum;
    for (int i = lo; i < hi; ++i) {
        out[i] = run;
        run += in[i];
    }
    if (threadIdx.x == T - 1) out[L] = s_part[T - 1];
}

__global__ void __launch_bounds__(1024)
k7_eval_scan_kernel(const int32_t* __restrict__ img_count, int B, int32_t* __restrict__ row_offsets,
                    const int32_t* __restrict__ class_count, int L, int32_t* __restrict__ class_offsets) {
    __shared__ int s_part[1024];
    k7_block_scan_array(img_count, row_offsets, B, s_part);
    if (class_offsets) k7_block_scan_array(class_count, class_offsets, L, s_part);
}

__device__ __forceinline__ void k7_eval_write_row(double* __restrict__ r, const double id, const float a, const float b2,
                                                  const float c, const float d, const float score, const double cat) {
    r[0] = id;
    r[1] = (double)a;
    r[2] = (double)b2;
    r[3] = (double)c;
    r[4] = (double)d;
    r[5] = (double)score;
    r[6] = cat;
}

// kernel 3: the rows.  Dynamic smem: cap ints (CSV: the label of every kept row, -1 otherwise).
__global__ void __launch_bounds__(K7_EVAL_THREADS)
k7_eval_write_kernel(const __grid_constant__ K7EvalDev E, const int32_t* __restrict__ row_offsets,
                     const int32_t* __restrict__ class_offsets, double* __restrict__ rows) {
    extern __shared__ int s_label[];
    __shared__ int s_w[K7_EVAL_THREADS / 32];
    __shared__ int s_min;
    const int b = blockIdx.x;
    const int n = max(0, min(E.counts[(size_t)b * E.counts_stride], E.cap));
    const float sc = E.scale[b];
    const float4* box4 = reinterpret_cast<const float4*>(E.boxes) + (size_t)b * E.cap;
    if (E.mode == K7_EVAL_COCO) {
        const int f = k7_eval_coco_stop(E, b, n, &s_min);
        const double id = E.image_ids ? (double)E.image_ids[b] : (double)b;
        int base = row_offsets[b];
        for (int i0 = 0; i0 < f; i0 += K7_EVAL_THREADS) {
            const int i = i0 + threadIdx.x;
            const size_t o = (size_t)b * E.cap + (i < f ? i : 0);
            const long long label = E.labels[o];
            const bool keep = i < f && k7_eval_label_ok(E, label);
            int total;
            const int r = k7_block_rank(keep, s_w, total);
            if (keep) {
                const float4 q = box4[i];
                // boxes /= scale (coco_eval.py:30), then w = x2 - x1, h = y2 - y1 on the scaled values (:34-35)
                const float x = __fdiv_rn(q.x, sc), y = __fdiv_rn(q.y, sc);
                const float w = __fsub_rn(__fdiv_rn(q.z, sc), x), h = __fsub_rn(__fdiv_rn(q.w, sc), y);
                const double cat = E.category_of_label ? (double)E.category_of_label[label] : (double)label;
                k7_eval_write_row(rows + (size_t)(base + r) * 7, id, x, y, w, h, E.scores[o], cat);
            }
            base += total;
        }
    } else {
        int emitted = 0;
        for (int i0 = 0; i0 < n; i0 += K7_EVAL_THREADS) {   // pass 1: which rows are kept
            const int i = i0 + threadIdx.x;
            const size_t o = (size_t)b * E.cap + (i < n ? i : 0);
            const long long label = E.labels[o];
            const bool sel = i < n && E.scores[o] > E.thr_f && k7_eval_label_ok(E, label);
            int total;
            const int r = emitted + k7_block_rank(sel, s_w, total);
            if (i < n) s_label[i] = (sel && (E.max_det <= 0 || r < E.max_det)) ? (int)label : -1;
            emitted += total;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += K7_EVAL_THREADS) {   // pass 2: rank within the label, in input order
            const int label = s_label[i];
            if (label < 0) continue;
            int r = 0;
            for (int j = 0; j < i; ++j) r += (s_label[j] == label);
            const float4 q = box4[i];
            const size_t pos = (size_t)class_offsets[(size_t)b * E.nc + label] + r;
            k7_eval_write_row(rows + pos * 7, (double)b, __fdiv_rn(q.x, sc), __fdiv_rn(q.y, sc), __fdiv_rn(q.z, sc),
                              __fdiv_rn(q.w, sc), E.scores[(size_t)b * E.cap + i], (double)label);
        }
    }
}

static void k7_eval_ws(int B, int nc, size_t* off_class, size_t* total) {
    size_t o = k7_align_up(sizeof(int32_t) * (size_t)B, 256);
    *off_class = o;
    o += k7_align_up(sizeof(int32_t) * (size_t)B * (size_t)(nc > 0 ? nc : 1), 256);
    *total = o;
}

extern "C" size_t k7_eval_pack_workspace_bytes(int B, int num_classes) {
    if (B <= 0 || num_classes < 0) return 0;
    size_t oc, total;
    k7_eval_ws(B, num_classes, &oc, &total);
    return total;
}

extern "C" int k7_eval_pack(const float* scores, const int64_t* labels, const float* boxes, int cap, const int32_t* counts,
                            int counts_stride, int B, const float* scale, const int64_t* image_ids,
                            const int32_t* category_of_label, const uint8_t* valid_label, const K7EvalParams* params,
                            double* rows, int32_t* row_offsets, int32_t* class_offsets, void* workspace,
                            size_t workspace_bytes, void* stream) {
    K7_REQUIRE(scores && labels && boxes && counts && scale && params && rows && row_offsets && workspace,
               "k7_eval_pack: null argument");
    K7_REQUIRE(B > 0 && cap > 0 && counts_stride > 0, "k7_eval_pack: bad shape B=%d cap=%d stride=%d", B, cap, counts_stride);
    K7_REQUIRE(params->mode == K7_EVAL_COCO || params->mode == K7_EVAL_CSV, "k7_eval_pack: bad mode %d", params->mode);
    K7_REQUIRE(params->num_classes > 0, "k7_eval_pack: num_classes must be positive");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15) == 0, "k7_eval_pack: boxes must be 16-byte aligned");
    const bool csv = params->mode == K7_EVAL_CSV;
    K7_REQUIRE(!csv || class_offsets, "k7_eval_pack: CSV mode needs class_offsets");
    K7_REQUIRE(!csv || (params->num_classes <= 8192 && cap <= 8192),
               "k7_eval_pack: CSV mode supports num_classes, cap <= 8192 (got %d, %d)", params->num_classes, cap);
    K7_REQUIRE((long long)B * cap < (1ll << 31) && (long long)B * params->num_classes < (1ll << 31), "k7_eval_pack: too large");
    size_t off_class, total;
    k7_eval_ws(B, params->num_classes, &off_class, &total);
    if (workspace_bytes < total) {
        k7_set_error("k7_eval_pack: workspace %zu < %zu bytes", workspace_bytes, total);
        return K7_EWORKSPACE;
    }
    unsigned char* ws = static_cast<unsigned char*>(workspace);
    int32_t* img_count = reinterpret_cast<int32_t*>(ws);
    int32_t* class_count = reinterpret_cast<int32_t*>(ws + off_class);
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    K7EvalDev E;
    E.scores = scores; E.labels = labels; E.boxes = boxes; E.counts = counts; E.scale = scale;
    E.image_ids = image_ids; E.category_of_label = category_of_label; E.valid_label = valid_label;
    E.cap = cap; E.counts_stride = counts_stride; E.B = B; E.mode = params->mode;
    E.max_det = params->max_detections; E.nc = params->num_classes;
    E.thr = params->score_thr;
    E.thr_f = (float)params->score_thr;   // csv_eval.py:99: numpy compares float32 scores with the float cast down

    k7_eval_count_kernel<<<B, K7_EVAL_THREADS, csv ? sizeof(int) * (size_t)E.nc : 0, st>>>(E, img_count, class_count);
    K7_LAUNCH_CHECK();
    k7_eval_scan_kernel<<<1, 1024, 0, st>>>(img_count, B, row_offsets, class_count, B * E.nc, csv ? class_offsets : nullptr);
    K7_LAUNCH_CHECK();
    k7_eval_write_kernel<<<B, K7_EVAL_THREADS, csv ? sizeof(int) * (size_t)cap : 0, st>>>(E, row_offsets, class_offsets, rows);
    K7_LAUNCH_CHECK();
    return K7_OK;
}
