// Anchor materialisation and anchor<->GT assignment (training side, stage 1).
//
//   k7_anchors_materialize   Anchors.forward           architectures/anchors.py:38-60
//   k7_match                 calc_iou + assignment     architectures/losses.py:6-23, 49-82
//
// k7_match is ALU-bound (A x M IoU pairs, no tensor larger than 4 B/anchor is written), so the
// design is about not doing pairs that cannot matter:
//   * the valid annotation rows of the image are compacted (order preserved) into shared memory
//     once per CTA, with their areas;
//   * each warp holds 32 consecutive anchors (a few neighbouring cells of one level); it reduces
//     their bounding hull with shuffles and then tests 32 GT boxes at a time against the hull,
//     one GT per lane; __ballot_sync gives the candidate set and only candidates are visited;
//   * a candidate whose clamped intersection with the lane's anchor is empty contributes IoU = +0
//     exactly, so the IEEE division is only executed for truly overlapping pairs.
// None of this changes a result bit: skipped pairs have IoU == 0, the running maximum starts at
// (0, index 0) and is only replaced on a strictly greater value, which is torch.max's
// first-index-on-ties rule (verified against the reference in tests/golden).
#include "k7_common.cuh"

// ------------------------------------------------------------------ anchors
__global__ void __launch_bounds__(256) k7_anchors_kernel(const __grid_constant__ K7AnchorSpec spec,
                                                         float4* __restrict__ out) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= spec.total) return;
    float x1, y1, x2, y2;
    k7_anchor_from_spec(spec, a, x1, y1, x2, y2);
    out[a] = (spec.box_order == K7_BOX_XYXY) ? make_float4(x1, y1, x2, y2) : make_float4(y1, x1, y2, x2);
}

extern "C" int k7_anchors_materialize(const K7AnchorSpec* spec, float* out, void* stream) {
    K7_REQUIRE(spec && out, "k7_anchors_materialize: null argument");
    int rc = k7_check_spec(spec, spec->total);
    if (rc) return rc;
    K7_REQUIRE((reinterpret_cast<uintptr_t>(out) & 15) == 0, "k7_anchors_materialize: out must be 16-byte aligned");
    const int threads = 256;
    const int blocks = (spec->total + threads - 1) / threads;
    k7_anchors_kernel<<<blocks, threads, 0, static_cast<cudaStream_t>(stream)>>>(*spec, reinterpret_cast<float4*>(out));
    K7_LAUNCH_CHECK();
    return K7_OK;
}

// ------------------------------------------------------------------ match
#define K7_MATCH_THREADS 256

template <bool FROM_SPEC>
__global__ void __launch_bounds__(K7_MATCH_THREADS)
k7_match_kernel(const __grid_constant__ K7AnchorSpec spec, const float* __restrict__ anchors, int box_order,
                const float* __restrict__ annotations, int A, int M, float neg_thr, float pos_thr,
                int32_t* __restrict__ assign, int32_t* __restrict__ num_pos, float* __restrict__ iou_max_out,
                int32_t* __restrict__ argmax_out, float4* __restrict__ zero_fill4) {
    extern __shared__ float smem[];
    // SoA of the valid GT rows: x1 | y1 | x2 | y2 | area | row (as int)
    float* gx1 = smem;
    float* gy1 = gx1 + M;
    float* gx2 = gy1 + M;
    float* gy2 = gx2 + M;
    float* gar = gy2 + M;
    int* grow = reinterpret_cast<int*>(gar + M);
    __shared__ int s_valid;
    __shared__ int s_pos[K7_MATCH_THREADS / 32];

    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float* ann = annotations + (size_t)b * M * 5;

    // order-preserving compaction of rows with label != -1 (losses.py:50), by warp 0
    if (warp == 0) {
        int base = 0;
        for (int m0 = 0; m0 < M; m0 += 32) {
            const int m = m0 + lane;
            float r0 = 0.f, r1 = 0.f, r2 = 0.f, r3 = 0.f, r4 = -1.f;
            if (m < M) {
                r0 = __ldg(ann + m * 5 + 0); r1 = __ldg(ann + m * 5 + 1); r2 = __ldg(ann + m * 5 + 2);
                r3 = __ldg(ann + m * 5 + 3); r4 = __ldg(ann + m * 5 + 4);
            }
            const bool ok = (m < M) && (r4 != -1.0f);
            const unsigned bal = __ballot_sync(0xffffffffu, ok);
            if (ok) {
                const int p = base + __popc(bal & ((1u << lane) - 1u));
                gx1[p] = r0; gy1[p] = r1; gx2[p] = r2; gy2[p] = r3;
                gar[p] = k7_area(r0, r1, r2, r3);
                grow[p] = m;
            }
            base += __popc(bal);
        }
        if (lane == 0) s_valid = base;
    }
    __syncthreads();
    const int nv = s_valid;

    const int a = blockIdx.x * K7_MATCH_THREADS + threadIdx.x;
    const bool live = a < A;
    float ax1 = 0.f, ay1 = 0.f, ax2 = 0.f, ay2 = 0.f;
    if (live) k7_anchor<FROM_SPEC>(spec, anchors, box_order, a, ax1, ay1, ax2, ay2);
    const float area_a = k7_area(ax1, ay1, ax2, ay2);

    float best = 0.0f;
    int besti = 0;
    if (nv > 0) {
        // hull of the warp's anchors (dead lanes contribute an empty hull)
        float hx1 = live ? ax1 : 3.0e38f, hy1 = live ? ay1 : 3.0e38f;
        float hx2 = live ? ax2 : -3.0e38f, hy2 = live ? ay2 : -3.0e38f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            hx1 = fminf(hx1, __shfl_xor_sync(0xffffffffu, hx1, o));
            hy1 = fminf(hy1, __shfl_xor_sync(0xffffffffu, hy1, o));
            hx2 = fmaxf(hx2, __shfl_xor_sync(0xffffffffu, hx2, o));
            hy2 = fmaxf(hy2, __shfl_xor_sync(0xffffffffu, hy2, o));
        }
        for (int g0 = 0; g0 < nv; g0 += 32) {
            const int g = g0 + lane;
            bool cand = false;
            if (g < nv) cand = (gx2[g] > hx1) && (gx1[g] < hx2) && (gy2[g] > hy1) && (gy1[g] < hy2);
            unsigned cm = __ballot_sync(0xffffffffu, cand);
            while (cm) {
                const int j = g0 + __ffs(cm) - 1;
                cm &= cm - 1;
                const float iou = k7_pair_iou(ax1, ay1, ax2, ay2, area_a, gx1[j], gy1[j], gx2[j], gy2[j], gar[j]);
                if (iou > best) {
                    best = iou;
                    besti = j;
                }
            }
        }
    }

    int state = K7_STATE_IGNORE;
    if (nv > 0) {
        if (best < neg_thr) state = K7_STATE_NEG;
        if (best >= pos_thr) state = K7_STATE_POS;
    }
    const bool pos = live && state == K7_STATE_POS;
    if (live) {
        const size_t o = (size_t)b * A + a;
        const int row = nv > 0 ? grow[besti] : 0;
        assign[o] = (row << 2) | state;
        if (iou_max_out) iou_max_out[o] = best;
        if (argmax_out) argmax_out[o] = nv > 0 ? besti : -1;
        if (zero_fill4) zero_fill4[o] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    const unsigned pb = __ballot_sync(0xffffffffu, pos);
    if (lane == 0) s_pos[warp] = __popc(pb);
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < K7_MATCH_THREADS / 32; ++w) t += s_pos[w];
        if (t) atomicAdd(num_pos + b, t);   // integer: order independent, exact
    }
}

extern "C" int k7_match(const K7AnchorSpec* spec, const float* anchors, int box_order, const float* annotations,
                        int B, int A, int M, const K7MatchParams* params, int32_t* assign, int32_t* num_pos,
                        float* iou_max, int32_t* argmax, float* zero_fill4, void* stream) {
    K7_REQUIRE(annotations && params && assign && num_pos, "k7_match: null argument");
    K7_REQUIRE(spec || anchors, "k7_match: need an anchor spec or an anchor tensor");
    K7_REQUIRE(B > 0 && A > 0 && M > 0, "k7_match: bad shape B=%d A=%d M=%d", B, A, M);
    K7_REQUIRE(M < (1 << 29), "k7_match: M too large");
    K7_REQUIRE(B <= 65535, "k7_match: B > 65535 not supported");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(zero_fill4) & 15) == 0, "k7_match: zero_fill4 must be 16-byte aligned");
    if (spec) {
        int rc = k7_check_spec(spec, A);
        if (rc) return rc;
    } else {
        K7_REQUIRE((reinterpret_cast<uintptr_t>(anchors) & 15) == 0, "k7_match: anchors must be 16-byte aligned");
        K7_REQUIRE(box_order == K7_BOX_XYXY || box_order == K7_BOX_YXYX, "k7_match: bad box_order");
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t smem = (size_t)M * 6 * sizeof(float);
    K7_REQUIRE(smem <= 200 * 1024, "k7_match: M=%d annotation rows exceed shared memory", M);
    K7_CUDA(cudaMemsetAsync(num_pos, 0, sizeof(int32_t) * B, st));
    dim3 grid((A + K7_MATCH_THREADS - 1) / K7_MATCH_THREADS, B);
    static const K7AnchorSpec kEmpty = {};
    if (spec) {
        if (smem > 48 * 1024)
            K7_CUDA(cudaFuncSetAttribute(k7_match_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k7_match_kernel<true><<<grid, K7_MATCH_THREADS, smem, st>>>(*spec, nullptr, spec->box_order, annotations, A, M,
                                                                    params->neg_thr, params->pos_thr, assign, num_pos,
                                                                    iou_max, argmax, reinterpret_cast<float4*>(zero_fill4));
    } else {
        if (smem > 48 * 1024)
            K7_CUDA(cudaFuncSetAttribute(k7_match_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        k7_match_kernel<false><<<grid, K7_MATCH_THREADS, smem, st>>>(kEmpty, anchors, box_order, annotations, A, M,
                                                                     params->neg_thr, params->pos_thr, assign, num_pos,
                                                                     iou_max, argmax, reinterpret_cast<float4*>(zero_fill4));
    }
    K7_LAUNCH_CHECK();
    return K7_OK;
}
