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
#include <stdlib.h>
#include <string.h>

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

// Host-computed helpers for the spec path: exact divisions by multiply-shift, and per level the
// extremes of the in-cell anchor shapes (used by the culling bound below).
struct K7MatchAux {
    K7FastDiv div_k;                      // / cell_anchors
    K7FastDiv div_fw[K7_MAX_LEVELS];      // / fw[level]
    float amin[K7_MAX_LEVELS], amax[K7_MAX_LEVELS];   // smallest / largest anchor area of the level
    float wmax[K7_MAX_LEVELS], hmax[K7_MAX_LEVELS];   // largest anchor width / height of the level
};

__device__ __forceinline__ int k7_level_of(const K7AnchorSpec& s, int a) {
    int l = 0;
#pragma unroll
    for (int i = 1; i < K7_MAX_LEVELS; ++i)
        if (i < s.num_levels && a >= s.offset[i]) l = i;
    return l;
}

// k7_anchor_from_spec with the three integer divisions done by multiply-shift
__device__ __forceinline__ void k7_anchor_fast(const K7AnchorSpec& s, const K7MatchAux& x, int a, int l, float& x1,
                                               float& y1, float& x2, float& y2) {
    const uint32_t r = (uint32_t)(a - s.offset[l]);
    const uint32_t cell = k7_div(r, x.div_k);
    const int k = (int)(r - cell * (uint32_t)s.cell_anchors);
    const uint32_t yy = k7_div(cell, x.div_fw[l]);
    const uint32_t xx = cell - yy * (uint32_t)s.fw[l];
    const double st = s.stride[l];
    const double sx = __dmul_rn(__dadd_rn((double)xx, 0.5), st);
    const double sy = __dmul_rn(__dadd_rn((double)yy, 0.5), st);
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

// One CTA = one image x a strided set of 256-anchor chunks, so the annotation rows are compacted
// once per CTA and reused.  CULL (production mode, no debug outputs): a GT can only change an
// anchor's state or a positive's arg-max if IoU >= neg_thr, which needs
//     neg_thr * area_a <= area_g <= area_a / neg_thr        and
//     inter_ub * (1 + neg_thr) >= neg_thr * (area_a + area_g),
// inter_ub = min(overlap_x(hull, g), wmax) * min(overlap_y(hull, g), hmax) >= inter(a, g) for every
// anchor a of the warp (a lies inside the hull).  Both are tested with a 1e-4 relative margin
// (float rounding is 1e-7), so no pair that reaches the threshold is ever skipped and the state /
// positive arg-max stay bit-exact; only the (unused) arg-max of negative anchors is unspecified.
template <bool FROM_SPEC, bool CULL>
__global__ void __launch_bounds__(K7_MATCH_THREADS)
k7_match_kernel(const __grid_constant__ K7AnchorSpec spec, const __grid_constant__ K7MatchAux aux,
                const float* __restrict__ anchors, int box_order, const float* __restrict__ annotations, int A, int M,
                float neg_thr, float pos_thr, int32_t* __restrict__ assign, int32_t* __restrict__ num_pos,
                float* __restrict__ iou_max_out, int32_t* __restrict__ argmax_out, float4* __restrict__ zero_fill4,
                uint32_t* __restrict__ ignore_bits, int W) {
    extern __shared__ float smem[];
    // valid GT rows: (x1,y1,x2,y2) as float4 | area | row (as int)
    float4* gbox = reinterpret_cast<float4*>(smem);
    float* gar = smem + 4 * M;
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
                gbox[p] = make_float4(r0, r1, r2, r3);
                gar[p] = k7_area(r0, r1, r2, r3);
                grow[p] = m;
            }
            base += __popc(bal);
        }
        if (lane == 0) s_valid = base;
    }
    __syncthreads();
    const int nv = s_valid;
    const float c_lo = neg_thr * (1.0f - 1e-4f);                      // area window / bound, with margin
    const float c_in = neg_thr / (1.0f + neg_thr) * (1.0f - 1e-4f);
    int pos_total = 0;

    for (int a0 = blockIdx.x * K7_MATCH_THREADS; a0 < A; a0 += gridDim.x * K7_MATCH_THREADS) {
        const int a = a0 + threadIdx.x;
        const bool live = a < A;
        float ax1 = 0.f, ay1 = 0.f, ax2 = 0.f, ay2 = 0.f;
        int lvl = 0;
        if (live) {
            if (FROM_SPEC) {
                lvl = k7_level_of(spec, a);
                k7_anchor_fast(spec, aux, a, lvl, ax1, ay1, ax2, ay2);
            } else {
                k7_anchor_from_tensor(anchors, box_order, a, ax1, ay1, ax2, ay2);
            }
        }
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
            float amin = 0.f, amax = 3.0e38f, wmax = 3.0e38f, hmax = 3.0e38f;
            if (CULL) {
                if (FROM_SPEC) {
                    // a warp's anchors lie in at most two adjacent levels: those of its first and last lane
                    const int l0 = __shfl_sync(0xffffffffu, lvl, 0);
                    const int l1 = k7_level_of(spec, min(a0 + (warp << 5) + 31, A - 1));
                    amin = fminf(aux.amin[l0], aux.amin[l1]);
                    amax = fmaxf(aux.amax[l0], aux.amax[l1]);
                    wmax = fmaxf(aux.wmax[l0], aux.wmax[l1]);
                    hmax = fmaxf(aux.hmax[l0], aux.hmax[l1]);
                } else {
                    amin = live ? area_a : 3.0e38f;
                    amax = live ? area_a : 0.0f;
                    wmax = live ? ax2 - ax1 : 0.0f;
                    hmax = live ? ay2 - ay1 : 0.0f;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        amin = fminf(amin, __shfl_xor_sync(0xffffffffu, amin, o));
                        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
                        wmax = fmaxf(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
                        hmax = fmaxf(hmax, __shfl_xor_sync(0xffffffffu, hmax, o));
                    }
                }
            }
            for (int g0 = 0; g0 < nv; g0 += 32) {
                const int g = g0 + lane;
                bool cand = false;
                if (g < nv) {
                    const float4 gb = gbox[g];
                    const float ow = fminf(hx2, gb.z) - fmaxf(hx1, gb.x), oh = fminf(hy2, gb.w) - fmaxf(hy1, gb.y);
                    cand = (ow > 0.0f) && (oh > 0.0f);
                    if (CULL) {
                        const float ag = gar[g];
                        const float ub = fminf(ow, wmax) * fminf(oh, hmax);
                        cand = cand && (ag >= c_lo * amin) && (ag * c_lo <= amax) && (ub >= c_in * (amin + ag));
                    }
                }
                unsigned cm = __ballot_sync(0xffffffffu, cand);
                while (cm) {
                    const int j = g0 + __ffs(cm) - 1;
                    cm &= cm - 1;
                    const float4 gb = gbox[j];
                    const float iou = k7_pair_iou(ax1, ay1, ax2, ay2, area_a, gb.x, gb.y, gb.z, gb.w, gar[j]);
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
        if (live) {
            const size_t o = (size_t)b * A + a;
            const int row = nv > 0 ? grow[besti] : 0;
            assign[o] = (row << 2) | state;
            if (iou_max_out) iou_max_out[o] = best;
            if (argmax_out) argmax_out[o] = nv > 0 ? besti : -1;
            if (zero_fill4) zero_fill4[o] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        pos_total += (live && state == K7_STATE_POS) ? 1 : 0;
        if (ignore_bits) {   // a0 is a multiple of 256: every warp owns whole words of the image's bitmap
            const unsigned ib = __ballot_sync(0xffffffffu, live && state == K7_STATE_IGNORE);
            if (lane == 0 && live) ignore_bits[(size_t)b * W + (a >> 5)] = ib;
        }
    }

    pos_total = k7_warp_sum(pos_total);
    if (lane == 0) s_pos[warp] = pos_total;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
#pragma unroll
        for (int w = 0; w < K7_MATCH_THREADS / 32; ++w) t += s_pos[w];
        if (t) atomicAdd(num_pos + b, t);   // integer: order independent, exact
    }
}

// ------------------------------------------------------------------ match, cell-mapped (anchors from a recipe)
// ncu on the anchor-mapped kernel above (C2): 400 instructions per anchor, of which the IoU tests
// are ~45 - the rest is per-THREAD fixed cost: anchor index -> (level, cell, shape), the double
// precision cell centre, the warp hull, the cull rounds over the ground truth.  All of that
// depends on the CELL, not on the anchor.  Here a lane owns a cell of an 8 x 4 tile of one
// pyramid level and loops over (a group of) the cell's anchor shapes: level / cell / centre /
// hull / candidate list once per tile, then per shape four double adds for the box and the IoU
// tests against the tile's candidates.  Results are bit-identical (same arithmetic per pair, same
// ascending visiting order -> same first-index ties); the hull and the area window are those of
// ONE level and one group of shapes, i.e. tighter than the anchor-mapped warp's.
// Work item = (level, tile, shape group); the groups split a tile when an image has too few tiles
// for the launch.  Ignore bits are OR-ed into a zeroed bitmap (about 1 % of the anchors).
#ifndef K7_MC_THREADS
#define K7_MC_THREADS 128   // warps of a CTA wait for its slowest one: small CTAs waste less
#endif
#define K7_MT_W 8
#define K7_MT_H 4
#define K7_MT_MAXG 4
struct K7MatchTiles {
    int item0[K7_MAX_LEVELS + 1];   // first work item of each level
    int tiles_x[K7_MAX_LEVELS];
    int groups, gsize;              // shape groups per tile, shapes per group
    double hb[K7_MAX_LEVELS][K7_MT_MAXG][4];   // hull of the group's shapes around the cell centre (x1, y1, x2, y2)
    float amin[K7_MAX_LEVELS][K7_MT_MAXG], amax[K7_MAX_LEVELS][K7_MT_MAXG];
    float wmax[K7_MAX_LEVELS][K7_MT_MAXG], hmax[K7_MAX_LEVELS][K7_MT_MAXG];
};

template <bool CULL>
__global__ void __launch_bounds__(K7_MC_THREADS)
k7_match_cells_kernel(const __grid_constant__ K7AnchorSpec spec, const __grid_constant__ K7MatchTiles T,
                      const float* __restrict__ annotations, int A, int M, float neg_thr, float pos_thr,
                      int32_t* __restrict__ assign, int32_t* __restrict__ num_pos, float* __restrict__ iou_max_out,
                      int32_t* __restrict__ argmax_out, float4* __restrict__ zero_fill4, uint32_t* __restrict__ ignore_bits,
                      int W) {
    extern __shared__ float smem[];
    float4* gbox = reinterpret_cast<float4*>(smem);
    float* gar = smem + 4 * M;
    int* grow = reinterpret_cast<int*>(gar + M);
    __shared__ int s_valid;
    __shared__ int s_pos[K7_MC_THREADS / 32];

    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int* cl = grow + M + warp * M;   // this warp's candidate list
    // with one shape group per tile, a tile row is ONE run of 8*K consecutive anchors: the words are
    // staged here and stored (with the zero fill) as full lines - lane-per-cell stores are 36 / 144
    // bytes apart and cost more than the kernel at EfficientDet-D7 (28 M anchors per batch)
    const int K = spec.cell_anchors;
    const bool staged = T.groups == 1;
    int* so = grow + M + (K7_MC_THREADS / 32) * M + warp * (K7_MT_H * K7_MT_W * K7_MAX_CELL_ANCHORS);
    const float* ann = annotations + (size_t)b * M * 5;

    // order-preserving compaction of rows with label != -1 (losses.py:50): every warp takes 32 rows per round
    // (one round for M <= 256: a single load latency instead of one per 32 rows)
    {
        __shared__ int s_cnt[K7_MC_THREADS / 32];
        int base = 0;
        for (int m0 = 0; m0 < M; m0 += K7_MC_THREADS) {
            const int m = m0 + threadIdx.x;
            float r0 = 0.f, r1 = 0.f, r2 = 0.f, r3 = 0.f, r4 = -1.f;
            if (m < M) {
                r0 = __ldg(ann + m * 5 + 0); r1 = __ldg(ann + m * 5 + 1); r2 = __ldg(ann + m * 5 + 2);
                r3 = __ldg(ann + m * 5 + 3); r4 = __ldg(ann + m * 5 + 4);
            }
            const bool ok = (m < M) && (r4 != -1.0f);
            const unsigned bal = __ballot_sync(0xffffffffu, ok);
            if (lane == 0) s_cnt[warp] = __popc(bal);
            __syncthreads();
            int before = base, total = 0;
#pragma unroll
            for (int w = 0; w < K7_MC_THREADS / 32; ++w) {
                const int c = s_cnt[w];
                if (w < warp) before += c;
                total += c;
            }
            if (ok) {
                const int p = before + __popc(bal & ((1u << lane) - 1u));
                gbox[p] = make_float4(r0, r1, r2, r3);
                gar[p] = k7_area(r0, r1, r2, r3);
                grow[p] = m;
            }
            base += total;
            __syncthreads();
        }
        if (threadIdx.x == 0) s_valid = base;
    }
    __syncthreads();
    const int nv = s_valid;
    if (nv == 0) {   // no annotation: every anchor is ignored (losses.py:52-60 skips the image)
        for (int a = blockIdx.x * K7_MC_THREADS + threadIdx.x; a < A; a += gridDim.x * K7_MC_THREADS) {
            const size_t o = (size_t)b * A + a;
            assign[o] = K7_STATE_IGNORE;
            if (iou_max_out) iou_max_out[o] = 0.0f;
            if (argmax_out) argmax_out[o] = -1;
            if (zero_fill4) zero_fill4[o] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (ignore_bits)
            for (int w = blockIdx.x * K7_MC_THREADS + threadIdx.x; w < W; w += gridDim.x * K7_MC_THREADS) {
                const int left = A - (w << 5);
                ignore_bits[(size_t)b * W + w] = left >= 32 ? 0xffffffffu : (left > 0 ? (1u << left) - 1u : 0u);
            }
        return;
    }
    const float c_lo = neg_thr * (1.0f - 1e-4f);                      // area window / bound, with margin
    const float c_in = neg_thr / (1.0f + neg_thr) * (1.0f - 1e-4f);
    const int total_items = T.item0[spec.num_levels];
    int pos_total = 0;

    // coarse levels first: their tiles meet the most ground truth and would otherwise be the tail
    for (int it = blockIdx.x * (K7_MC_THREADS / 32) + warp; it < total_items;
         it += gridDim.x * (K7_MC_THREADS / 32)) {
        const int item = total_items - 1 - it;
        int l = 0;
        while (l + 1 < spec.num_levels && item >= T.item0[l + 1]) ++l;   // warp-uniform
        int t = item - T.item0[l];
        const int grp = t % T.groups;
        t /= T.groups;
        const int tyi = t / T.tiles_x[l], txi = t - tyi * T.tiles_x[l];
        const int fw = spec.fw[l], fh = spec.fh[l];
        const int x = txi * K7_MT_W + (lane & (K7_MT_W - 1)), y = tyi * K7_MT_H + (lane / K7_MT_W);
        const bool live = x < fw && y < fh;
        const double st = spec.stride[l];
        const double sx = __dmul_rn(__dadd_rn((double)x, 0.5), st);
        const double sy = __dmul_rn(__dadd_rn((double)y, 0.5), st);
        // hull of the tile's anchors of this group (float rounding is monotone: the hull of the
        // rounded extremes contains every rounded anchor)
        float hx1 = live ? __double2float_rn(__dadd_rn(T.hb[l][grp][0], sx)) : 3.0e38f;
        float hy1 = live ? __double2float_rn(__dadd_rn(T.hb[l][grp][1], sy)) : 3.0e38f;
        float hx2 = live ? __double2float_rn(__dadd_rn(T.hb[l][grp][2], sx)) : -3.0e38f;
        float hy2 = live ? __double2float_rn(__dadd_rn(T.hb[l][grp][3], sy)) : -3.0e38f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            hx1 = fminf(hx1, __shfl_xor_sync(0xffffffffu, hx1, o));
            hy1 = fminf(hy1, __shfl_xor_sync(0xffffffffu, hy1, o));
            hx2 = fmaxf(hx2, __shfl_xor_sync(0xffffffffu, hx2, o));
            hy2 = fmaxf(hy2, __shfl_xor_sync(0xffffffffu, hy2, o));
        }
        const float amin = T.amin[l][grp], amax = T.amax[l][grp], wmax = T.wmax[l][grp], hmax = T.hmax[l][grp];
        // candidate ground truth of the tile, ascending
        int nc = 0;
        for (int g0 = 0; g0 < nv; g0 += 32) {
            const int g = g0 + lane;
            bool cand = false;
            if (g < nv) {
                const float4 gb = gbox[g];
                const float ow = fminf(hx2, gb.z) - fmaxf(hx1, gb.x), oh = fminf(hy2, gb.w) - fmaxf(hy1, gb.y);
                cand = (ow > 0.0f) && (oh > 0.0f);
                if (CULL) {
                    const float ag = gar[g];
                    const float ub = fminf(ow, wmax) * fminf(oh, hmax);
                    cand = cand && (ag >= c_lo * amin) && (ag * c_lo <= amax) && (ub >= c_in * (amin + ag));
                }
            }
            const unsigned cm = __ballot_sync(0xffffffffu, cand);
            if (cand) cl[nc + __popc(cm & ((1u << lane) - 1u))] = g;
            nc += __popc(cm);
        }
        __syncwarp();
        const int k0 = grp * T.gsize, k1 = min(k0 + T.gsize, K);
        for (int k = k0; k < k1; ++k) {
            const double* bb = spec.base[l][k];
            float ax1, ay1, ax2, ay2;
            if (spec.box_order == K7_BOX_XYXY) {
                ax1 = __double2float_rn(__dadd_rn(bb[0], sx));
                ay1 = __double2float_rn(__dadd_rn(bb[1], sy));
                ax2 = __double2float_rn(__dadd_rn(bb[2], sx));
                ay2 = __double2float_rn(__dadd_rn(bb[3], sy));
            } else {
                ay1 = __double2float_rn(__dadd_rn(bb[0], sy));
                ax1 = __double2float_rn(__dadd_rn(bb[1], sx));
                ay2 = __double2float_rn(__dadd_rn(bb[2], sy));
                ax2 = __double2float_rn(__dadd_rn(bb[3], sx));
            }
            const float area_a = k7_area(ax1, ay1, ax2, ay2);
            float best = 0.0f;
            int besti = 0;
            for (int c = 0; c < nc; ++c) {
                const int j = cl[c];
                const float4 gb = gbox[j];
                const float iou = k7_pair_iou(ax1, ay1, ax2, ay2, area_a, gb.x, gb.y, gb.z, gb.w, gar[j]);
                if (iou > best) {
                    best = iou;
                    besti = j;
                }
            }
            int state = K7_STATE_IGNORE;
            if (best < neg_thr) state = K7_STATE_NEG;
            if (best >= pos_thr) state = K7_STATE_POS;
            if (live) {
                const int a = spec.offset[l] + (y * fw + x) * K + k;
                const size_t o = (size_t)b * A + a;
                const int word = (grow[besti] << 2) | state;
                if (staged) {
                    so[((lane / K7_MT_W) * K7_MT_W + (lane & (K7_MT_W - 1))) * K + k] = word;
                } else {
                    assign[o] = word;
                    if (zero_fill4) zero_fill4[o] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
                if (iou_max_out) iou_max_out[o] = best;
                if (argmax_out) argmax_out[o] = besti;
                if (ignore_bits && state == K7_STATE_IGNORE) atomicOr(ignore_bits + (size_t)b * W + (a >> 5), 1u << (a & 31));
                pos_total += state == K7_STATE_POS ? 1 : 0;
            }
        }
        __syncwarp();   // the candidate list is rewritten by the next item
        if (staged) {
            const int ncell = min(K7_MT_W, fw - txi * K7_MT_W), run = ncell * K;
            for (int r = 0; r < K7_MT_H; ++r) {
                const int yr = tyi * K7_MT_H + r;
                if (yr >= fh) break;
                const size_t o0 = (size_t)b * A + spec.offset[l] + (size_t)(yr * fw + txi * K7_MT_W) * K;
                for (int i = lane; i < run; i += 32) {
                    assign[o0 + i] = so[r * K7_MT_W * K + i];
                    if (zero_fill4) zero_fill4[o0 + i] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            __syncwarp();
        }
    }

    pos_total = k7_warp_sum(pos_total);
    if (lane == 0) s_pos[warp] = pos_total;
    __syncthreads();
    if (threadIdx.x == 0) {
        int tsum = 0;
#pragma unroll
        for (int w = 0; w < K7_MC_THREADS / 32; ++w) tsum += s_pos[w];
        if (tsum) atomicAdd(num_pos + b, tsum);   // integer: order independent, exact
    }
}

// work items and per-(level, group) culling extremes of the cell-mapped kernel
static void k7_match_tiles(const K7AnchorSpec* s, int B, K7MatchTiles* T) {
    memset(T, 0, sizeof(*T));
    const int K = s->cell_anchors;
    int tiles = 0;
    for (int l = 0; l < s->num_levels; ++l)
        tiles += ((s->fw[l] + K7_MT_W - 1) / K7_MT_W) * ((s->fh[l] + K7_MT_H - 1) / K7_MT_H);
    // enough items for three per resident warp of the image's share of the launch
    const long long warps_per_image = (8ll * (256 / K7_MC_THREADS) * k7_sm_count() * (K7_MC_THREADS / 32) + B - 1) / B;
    int groups = 1;
    while (groups < K && groups < K7_MT_MAXG && (long long)tiles * groups < 3 * warps_per_image) ++groups;
    T->gsize = (K + groups - 1) / groups;
    T->groups = (K + T->gsize - 1) / T->gsize;
    int item = 0;
    for (int l = 0; l < s->num_levels; ++l) {
        T->item0[l] = item;
        T->tiles_x[l] = (s->fw[l] + K7_MT_W - 1) / K7_MT_W;
        item += T->tiles_x[l] * ((s->fh[l] + K7_MT_H - 1) / K7_MT_H) * T->groups;
        for (int g = 0; g < T->groups; ++g) {
            double hx1 = 1e300, hy1 = 1e300, hx2 = -1e300, hy2 = -1e300, amin = 1e300, amax = 0, wmax = 0, hmax = 0;
            for (int k = g * T->gsize; k < K && k < (g + 1) * T->gsize; ++k) {
                const double* bq = s->base[l][k];
                const bool xy = s->box_order == K7_BOX_XYXY;
                const double x1 = xy ? bq[0] : bq[1], y1 = xy ? bq[1] : bq[0], x2 = xy ? bq[2] : bq[3], y2 = xy ? bq[3] : bq[2];
                const double w = x2 - x1, h = y2 - y1, ar = w * h;
                if (x1 < hx1) hx1 = x1;
                if (y1 < hy1) hy1 = y1;
                if (x2 > hx2) hx2 = x2;
                if (y2 > hy2) hy2 = y2;
                if (ar < amin) amin = ar;
                if (ar > amax) amax = ar;
                if (w > wmax) wmax = w;
                if (h > hmax) hmax = h;
            }
            T->hb[l][g][0] = hx1; T->hb[l][g][1] = hy1; T->hb[l][g][2] = hx2; T->hb[l][g][3] = hy2;
            // outward rounding: the device compares float areas of float-rounded anchors against these
            T->amin[l][g] = (float)(amin * (1.0 - 1e-5));
            T->amax[l][g] = (float)(amax * (1.0 + 1e-5));
            T->wmax[l][g] = (float)(wmax * (1.0 + 1e-5) + 1e-3);
            T->hmax[l][g] = (float)(hmax * (1.0 + 1e-5) + 1e-3);
        }
    }
    for (int l = s->num_levels; l <= K7_MAX_LEVELS; ++l) T->item0[l] = item;
}

static void k7_match_aux(const K7AnchorSpec* s, K7MatchAux* x) {
    memset(x, 0, sizeof(*x));
    x->div_k = k7_fastdiv_make((uint32_t)s->cell_anchors);
    for (int l = 0; l < s->num_levels; ++l) {
        x->div_fw[l] = k7_fastdiv_make((uint32_t)(s->fw[l] > 0 ? s->fw[l] : 1));
        double amin = 1e300, amax = 0, wmax = 0, hmax = 0;
        for (int k = 0; k < s->cell_anchors; ++k) {
            const double* b = s->base[l][k];
            // extents are the same for both storage orders up to naming; take both and keep the extremes
            const double d0 = b[2] - b[0], d1 = b[3] - b[1];
            const double w = s->box_order == K7_BOX_XYXY ? d0 : d1, h = s->box_order == K7_BOX_XYXY ? d1 : d0;
            const double ar = w * h;
            if (ar < amin) amin = ar;
            if (ar > amax) amax = ar;
            if (w > wmax) wmax = w;
            if (h > hmax) hmax = h;
        }
        // outward rounding: the device compares float areas of float-rounded anchors against these
        x->amin[l] = (float)(amin * (1.0 - 1e-5));
        x->amax[l] = (float)(amax * (1.0 + 1e-5));
        x->wmax[l] = (float)(wmax * (1.0 + 1e-5) + 1e-3);
        x->hmax[l] = (float)(hmax * (1.0 + 1e-5) + 1e-3);
    }
}

extern "C" int k7_match(const K7AnchorSpec* spec, const float* anchors, int box_order, const float* annotations,
                        int B, int A, int M, const K7MatchParams* params, int32_t* assign, int32_t* num_pos,
                        float* iou_max, int32_t* argmax, float* zero_fill4, uint32_t* ignore_bits, void* stream) {
    K7_REQUIRE(annotations && params && assign && num_pos, "k7_match: null argument");
    K7_REQUIRE(spec || anchors, "k7_match: need an anchor spec or an anchor tensor");
    K7_REQUIRE(B > 0 && A > 0 && M > 0, "k7_match: bad shape B=%d A=%d M=%d", B, A, M);
    K7_REQUIRE(M < (1 << 29), "k7_match: M too large");
    K7_REQUIRE(B <= 65535, "k7_match: B > 65535 not supported");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(zero_fill4) & 15) == 0, "k7_match: zero_fill4 must be 16-byte aligned");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(ignore_bits) & 15) == 0, "k7_match: ignore_bits must be 16-byte aligned");
    K7MatchAux aux;
    memset(&aux, 0, sizeof(aux));
    if (spec) {
        int rc = k7_check_spec(spec, A);
        if (rc) return rc;
        k7_match_aux(spec, &aux);
        box_order = spec->box_order;
    } else {
        K7_REQUIRE((reinterpret_cast<uintptr_t>(anchors) & 15) == 0, "k7_match: anchors must be 16-byte aligned");
        K7_REQUIRE(box_order == K7_BOX_XYXY || box_order == K7_BOX_YXYX, "k7_match: bad box_order");
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t smem = (size_t)M * 6 * sizeof(float);
    K7_REQUIRE(smem <= 200 * 1024, "k7_match: M=%d annotation rows exceed shared memory", M);
    K7_CUDA(cudaMemsetAsync(num_pos, 0, sizeof(int32_t) * B, st));
    const int chunks = (A + K7_MATCH_THREADS - 1) / K7_MATCH_THREADS;
    int gx = (8 * k7_sm_count() + B - 1) / B;   // about eight CTAs per SM over the batch
    if (gx > chunks) gx = chunks;
    if (gx < 1) gx = 1;
    dim3 grid(gx, B);
    // the exact arg-max of every anchor is only needed for the debug outputs; thresholds outside (0,1) disable culling
    const bool cull = !iou_max && !argmax && params->neg_thr > 1e-3f && params->neg_thr < 1.0f &&
                      params->pos_thr >= params->neg_thr;
    static const bool cells_on = [] { const char* e = getenv("K7_MATCH_CELLS"); return !e || atoi(e) != 0; }();
    if (spec && cells_on && (size_t)M * (6 + K7_MC_THREADS / 32) * sizeof(float) <= 160 * 1024) {   // anchors from a recipe: the cell-mapped kernel
        K7MatchTiles T;
        k7_match_tiles(spec, B, &T);
        const size_t smem_c = ((size_t)M * (6 + K7_MC_THREADS / 32) +
                               (K7_MC_THREADS / 32) * K7_MT_H * K7_MT_W * K7_MAX_CELL_ANCHORS) * sizeof(float);
        const int W = (int)K7_IGNORE_WORDS(A);
        if (ignore_bits) K7_CUDA(cudaMemsetAsync(ignore_bits, 0, sizeof(uint32_t) * (size_t)B * W, st));
        const int items = T.item0[spec->num_levels];
        int gxc = (items + K7_MC_THREADS / 32 - 1) / (K7_MC_THREADS / 32);
        const int gx_cap = (8 * (256 / K7_MC_THREADS) * k7_sm_count() + B - 1) / B;
        if (gxc > gx_cap) gxc = gx_cap;
        if (gxc < 1) gxc = 1;
        auto kern = cull ? k7_match_cells_kernel<true> : k7_match_cells_kernel<false>;
        if (smem_c > 48 * 1024) K7_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_c));
        kern<<<dim3(gxc, B), K7_MC_THREADS, smem_c, st>>>(*spec, T, annotations, A, M, params->neg_thr, params->pos_thr,
                                                              assign, num_pos, iou_max, argmax,
                                                              reinterpret_cast<float4*>(zero_fill4), ignore_bits, W);
        K7_LAUNCH_CHECK();
        return K7_OK;
    }
    static const K7AnchorSpec kEmpty = {};
    const K7AnchorSpec& sp = spec ? *spec : kEmpty;
#define K7_MATCH_LAUNCH(SP, CU)                                                                                          \
    do {                                                                                                                 \
        auto kern = k7_match_kernel<SP, CU>;                                                                             \
        if (smem > 48 * 1024)                                                                                            \
            K7_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));                 \
        kern<<<grid, K7_MATCH_THREADS, smem, st>>>(sp, aux, anchors, box_order, annotations, A, M, params->neg_thr,      \
                                                   params->pos_thr, assign, num_pos, iou_max, argmax,                    \
                                                   reinterpret_cast<float4*>(zero_fill4), ignore_bits,                   \
                                                   (int)K7_IGNORE_WORDS(A));                                             \
    } while (0)
    if (spec) { if (cull) K7_MATCH_LAUNCH(true, true); else K7_MATCH_LAUNCH(true, false); }
    else      { if (cull) K7_MATCH_LAUNCH(false, true); else K7_MATCH_LAUNCH(false, false); }
#undef K7_MATCH_LAUNCH
    K7_LAUNCH_CHECK();
    return K7_OK;
}

// ------------------------------------------------------------------ match, ground-truth-first (the production path)
// The two kernels above visit every anchor and, per tile, every ground-truth box that survives a
// cull: A x (a few) IoU tests plus a fixed cost per anchor - 51 us at EfficientDet-D0 (3.1 M
// anchors), although only ~2 % of the anchors end up in a state other than NEGATIVE.  Here the
// loop is turned round.  A pair can only matter when IoU >= neg_thr, and
//     IoU >= t   <=>   inter * (1 + t) >= t * (area_a + area_g)
// bounds, for one ground-truth box g and one anchor SHAPE (level l, in-cell anchor k: all anchors
// of a shape are translates of each other),
//     the areas:          t * area_a <= area_g <= area_a / t                 (IoU <= area ratio)
//     the intersection:   iw >= imin / min(h_a, h_g),  ih >= imin / min(w_a, w_g),
//                         imin = t/(1+t) * (area_a + area_g)                 (iw <= min(w), ih <= min(h))
// and iw >= iw_min is an interval of cell columns, ih >= ih_min an interval of cell rows.  One warp
// takes (image, annotation row), works out those rectangles for all (level, shape) pairs at once and
// evaluates the reference's exact float32 IoU (k7_pair_iou) on their cells - typically 5 x 5 per
// shape that passes the area window.  Every bound is taken with a 1e-4 relative margin and the intervals are
// widened by 0.02 cell, against float32 roundings of ~1e-7: no pair that reaches neg_thr is missed,
// so states, num_pos and the arg-max of every non-negative anchor are those of the exhaustive
// kernels, bit for bit (tests/test_gpu_parity.py).  Pairs at or above neg_thr are folded into
// best[] with a 64-bit atomicMax: generation | IoU bits | ~row - the largest IoU, then the lowest
// row (torch.max returns the first index).  The maximum is order independent.  The value the atomic
// returns tells whether the entry was written for the first time in this generation: those anchors
// (~2-3 % of A) are appended to the TOUCHED list behind the array, which is all k7_loss reads - no
// pass over the B*A entries anywhere.
struct __align__(16) K7ScatterShape {   // one anchor shape (level l, in-cell anchor k), 80 bytes
    double base[4];                   // the recipe's in-cell box (spec.base[l][k])
    float g_lo, g_hi;                 // window of ground-truth areas that can reach neg_thr with this shape
    float area_lo;                    // its area, rounded down
    float wa, ha;                     // its extents, rounded up
    float bx1, bx2, by1, by2;         // its box around the cell centre as (x1, x2, y1, y2)
    float pad[3];
};
struct K7ScatterAux {
    float lvl_lo[K7_MAX_LEVELS], lvl_hi[K7_MAX_LEVELS];   // union of the windows of a level's shapes
    float inv_stride[K7_MAX_LEVELS];
    float c_in;                                           // t/(1+t) * (1 - 1e-4)
    int pad[3];
    K7ScatterShape shapes[K7_MAX_LEVELS * K7_MAX_CELL_ANCHORS];   // [l * cell_anchors + k]
};

#define K7_SCATTER_THREADS 128
#define K7_SCATTER_U 1             // candidates per lane and trip (4 measured: no faster - the kernel is not bound by the atomics round trips)
#define K7_SCATTER_PAIRS (K7_MAX_LEVELS * K7_MAX_CELL_ANCHORS)

struct K7ScatterLevel {   // per pyramid level, in shared memory
    double stride;
    float stride_f;
    float inv_stride, lvl_lo, lvl_hi;
    int fw, fh, offset;
};

// Work item = (image, annotation row), one warp.  Lane p (two rounds when levels * shapes > 32) looks at
// anchor shape p = level * K + k and works out the rectangle of cells whose anchors of that shape can reach
// neg_thr; a warp scan turns the rectangle sizes into offsets, and then ALL candidate cells of ALL shapes
// are one flat range that the 32 lanes walk together (a box has ~50-150 candidates on 2-3 levels: two to
// five trips instead of one trip per shape with most lanes idle).  Every warp - also those without a work item -
// first - zero the d(reg) buffer of the loss stage (16 B per anchor: the kernel is latency bound, the stores
// drain while the work items run).
__global__ void __launch_bounds__(K7_SCATTER_THREADS)
k7_match_scatter_kernel(const __grid_constant__ K7AnchorSpec spec, const __grid_constant__ K7ScatterAux X,
                        const float* __restrict__ annotations, int B, int A, int M, float neg_thr,
                        unsigned long long* __restrict__ best, unsigned int gen, unsigned int* __restrict__ touched_count,
                        unsigned int* __restrict__ touched, float4* __restrict__ zero_fill4) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int L = spec.num_levels, K = spec.cell_anchors, LK = L * K;
    K7ScatterShape* sh = reinterpret_cast<K7ScatterShape*>(smem_raw);
    K7ScatterLevel* lv = reinterpret_cast<K7ScatterLevel*>(smem_raw + sizeof(K7ScatterShape) * (size_t)LK);
    __shared__ int s_pref[K7_SCATTER_THREADS / 32][K7_SCATTER_PAIRS];        // inclusive prefix of the candidate counts
    __shared__ uint2 s_rect[K7_SCATTER_THREADS / 32][K7_SCATTER_PAIRS];      // x: xlo | ylo << 16, y: cells per row
    {
        const int n16 = LK * (int)(sizeof(K7ScatterShape) / 16);
        const uint4* src = reinterpret_cast<const uint4*>(X.shapes);
        uint4* dst = reinterpret_cast<uint4*>(sh);
        for (int i = threadIdx.x; i < n16; i += K7_SCATTER_THREADS) dst[i] = src[i];
        if ((int)threadIdx.x < L) {
            const int l = threadIdx.x;
            K7ScatterLevel v;
            v.stride = spec.stride[l];
            v.stride_f = (float)spec.stride[l];
            v.inv_stride = X.inv_stride[l];
            v.lvl_lo = X.lvl_lo[l];
            v.lvl_hi = X.lvl_hi[l];
            v.fw = spec.fw[l];
            v.fh = spec.fh[l];
            v.offset = spec.offset[l];
            lv[l] = v;
        }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const long long gwarp = ((long long)blockIdx.x * K7_SCATTER_THREADS + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * K7_SCATTER_THREADS) >> 5;
    const long long tasks = (long long)B * M;
    const unsigned long long tag = (unsigned long long)gen << 48;
    const bool xy = spec.box_order == K7_BOX_XYXY;
    int* pref = s_pref[wib];
    uint2* rect = s_rect[wib];
    // the zero fill first: the stores drain while the (latency-bound) work items below run
    if (zero_fill4) {
        const size_t n = (size_t)B * (size_t)A;
        const size_t step = (size_t)gridDim.x * K7_SCATTER_THREADS;
        const float4 z = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        for (size_t i = (size_t)blockIdx.x * K7_SCATTER_THREADS + threadIdx.x; i < n; i += step) __stcs(zero_fill4 + i, z);
    }
    for (long long t = gwarp; t < tasks; t += nwarps) {
        const int m = (int)(t % M), b = (int)(t / M);
        const float* row = annotations + (size_t)t * 5;
        float gv = lane < 5 ? __ldg(row + lane) : 0.0f;
        const float g0 = __shfl_sync(0xffffffffu, gv, 0), g1 = __shfl_sync(0xffffffffu, gv, 1);
        const float g2 = __shfl_sync(0xffffffffu, gv, 2), g3 = __shfl_sync(0xffffffffu, gv, 3);
        const float g4 = __shfl_sync(0xffffffffu, gv, 4);
        if (g4 == -1.0f) continue;                        // pad row (losses.py:50)
        const float gw = K7_FSUB(g2, g0), gh = K7_FSUB(g3, g1);
        if (!(gw > 0.0f && gh > 0.0f)) continue;          // empty box: IoU is 0 with every anchor
        const float area_g = k7_area(g0, g1, g2, g3);
        const bool fast_gt = gw >= 8.0f && gh >= 8.0f;
        const float thr_lo = neg_thr - 1e-3f;
        int total = 0;
        __syncwarp();                                     // the previous item's tables are no longer read
        for (int p0 = 0; p0 < LK; p0 += 32) {             // warp-uniform
            const int p = p0 + lane;
            int n = 0, nx = 0, xlo = 0, ylo = 0;
            if (p < LK) {
                const int l = p / K;
                const K7ScatterLevel& V = lv[l];
                const K7ScatterShape& S = sh[p];
                if (V.fw > 0 && V.fh > 0 && area_g >= S.g_lo && area_g <= S.g_hi) {
                    const float imin = X.c_in * (S.area_lo + area_g);
                    const float mw = fminf(S.wa, gw), mh = fminf(S.ha, gh);
                    if (!(imin > mw * mh)) {
                        const float iw_min = imin / mh, ih_min = imin / mw;
                        const float inv_s = V.inv_stride;
                        // cell centre c = (x + 0.5) * stride must satisfy  g0 + iw_min - bx2 <= c <= g2 - iw_min - bx1
                        const float xl = (g0 + iw_min - S.bx2) * inv_s - 0.52f, xh = (g2 - iw_min - S.bx1) * inv_s - 0.48f;
                        const float yl = (g1 + ih_min - S.by2) * inv_s - 0.52f, yh = (g3 - ih_min - S.by1) * inv_s - 0.48f;
                        xlo = max(0, (int)ceilf(fmaxf(xl, -1.0f)));
                        ylo = max(0, (int)ceilf(fmaxf(yl, -1.0f)));
                        const int xhi = min(V.fw - 1, (int)floorf(fminf(xh, (float)V.fw)));
                        const int yhi = min(V.fh - 1, (int)floorf(fminf(yh, (float)V.fh)));
                        if (xlo <= xhi && ylo <= yhi) {
                            nx = xhi - xlo + 1;
                            n = nx * (yhi - ylo + 1);
                        }
                    }
                }
            }
            int inc = n;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += v;
            }
            if (p < LK) {
                pref[p] = total + inc;
                rect[p] = make_uint2((unsigned)xlo | ((unsigned)ylo << 16), (unsigned)nx);
            }
            total += __shfl_sync(0xffffffffu, inc, 31);
        }
        __syncwarp();
        // K7_SCATTER_U candidates per lane and trip: their atomics are issued back to back and answered together, and
        // the list takes ONE reservation per trip.
        constexpr int U = K7_SCATTER_U;
        for (int i0 = 0; i0 < total; i0 += 32 * U) {      // warp-uniform
            int an[U];
            unsigned long long val[U];
            bool hit[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = i0 + 32 * u + lane;
                hit[u] = false;
                an[u] = 0;
                val[u] = 0ull;
                if (i >= total) continue;
                int lo = 0, hi = LK - 1;                  // the first shape whose inclusive prefix exceeds i
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    if (pref[mid] > i) hi = mid; else lo = mid + 1;
                }
                const int p = lo, l = p / K, k = p - l * K;
                const uint2 rc = rect[p];
                const int nx = (int)rc.y, j = i - (p ? pref[p - 1] : 0);
                const int yy = j / nx, x = (int)(rc.x & 0xffffu) + (j - yy * nx), y = (int)(rc.x >> 16) + yy;
                const K7ScatterLevel& V = lv[l];
                const K7ScatterShape& S = sh[p];
                // A cheap float32 look first (most candidates of the superset fail): the anchor from float32 parts is
                // within 1e-4 px of the reference's, so for extents >= 8 px its IoU is within 1e-4 of the exact one;
                // pairs below neg_thr - 1e-3 leave here, division-free.  The exact float64 / reference-order path
                // below decides everything that stays.
                if (fast_gt && S.wa >= 8.0f && S.ha >= 8.0f) {
                    const float cxf = ((float)x + 0.5f) * V.stride_f, cyf = ((float)y + 0.5f) * V.stride_f;
                    const float fw_ = fminf(cxf + S.bx2, g2) - fmaxf(cxf + S.bx1, g0);
                    const float fh_ = fminf(cyf + S.by2, g3) - fmaxf(cyf + S.by1, g1);
                    const float fi = fw_ * fh_;
                    const float fu = (S.bx2 - S.bx1) * (S.by2 - S.by1) + area_g - fi;
                    if (!(fw_ > 0.0f && fh_ > 0.0f) || fi < thr_lo * fu) continue;
                }
                const double sx = __dmul_rn(__dadd_rn((double)x, 0.5), V.stride);
                const double sy = __dmul_rn(__dadd_rn((double)y, 0.5), V.stride);
                float ax1, ay1, ax2, ay2;                 // k7_anchor_from_spec: float64 add, one rounding
                if (xy) {
                    ax1 = __double2float_rn(__dadd_rn(S.base[0], sx));
                    ay1 = __double2float_rn(__dadd_rn(S.base[1], sy));
                    ax2 = __double2float_rn(__dadd_rn(S.base[2], sx));
                    ay2 = __double2float_rn(__dadd_rn(S.base[3], sy));
                } else {
                    ay1 = __double2float_rn(__dadd_rn(S.base[0], sy));
                    ax1 = __double2float_rn(__dadd_rn(S.base[1], sx));
                    ay2 = __double2float_rn(__dadd_rn(S.base[2], sy));
                    ax2 = __double2float_rn(__dadd_rn(S.base[3], sx));
                }
                const float iou = k7_pair_iou(ax1, ay1, ax2, ay2, k7_area(ax1, ay1, ax2, ay2), g0, g1, g2, g3, area_g);
                if (iou >= neg_thr) {
                    hit[u] = true;
                    an[u] = V.offset + (y * V.fw + x) * K + k;
                    val[u] = tag | ((unsigned long long)__float_as_uint(iou) << 16) | (unsigned long long)(0xFFFF - m);
                }
            }
            unsigned long long old[U];
#pragma unroll
            for (int u = 0; u < U; ++u) old[u] = hit[u] ? atomicMax(best + (size_t)b * A + an[u], val[u]) : ~0ull;
            unsigned int fm[U];
            int nfirst = 0;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                // nobody wrote this entry in this generation before (a lane that did not hit reads ~0: never first)
                fm[u] = __ballot_sync(0xffffffffu, hit[u] && K7_BEST_GEN(old[u]) != gen);
                nfirst += __popc(fm[u]);
            }
            if (nfirst == 0) continue;                    // warp-uniform
            unsigned int base = 0u;
            if (lane == 0) base = atomicAdd(touched_count + (size_t)b * K7_BEST_CNT_STRIDE, (unsigned int)nfirst);
            base = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if ((fm[u] >> lane) & 1u) touched[(size_t)b * A + base + (unsigned int)__popc(fm[u] & ((1u << lane) - 1u))] = (unsigned int)an[u];
                base += (unsigned int)__popc(fm[u]);
            }
        }
    }
}

extern "C" size_t k7_match_scatter_workspace_bytes(int B, int A) {
    if (B <= 0 || A <= 0) return 0;
    return k7_best_list_off(B, A) + k7_best_list_hdr(B) + sizeof(unsigned int) * (size_t)B * (size_t)A;
}

extern "C" int k7_match_scatter(const K7AnchorSpec* spec, const float* annotations, int B, int A, int M,
                                const K7MatchParams* params, uint64_t* best, uint32_t generation, float* zero_fill4,
                                void* stream) {
    K7_REQUIRE(spec && annotations && params && best, "k7_match_scatter: null argument");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(zero_fill4) & 15) == 0, "k7_match_scatter: zero_fill4 must be 16-byte aligned");
    K7_REQUIRE(B > 0 && B <= 65535 && A > 0 && M > 0 && M <= 65535, "k7_match_scatter: bad shape B=%d A=%d M=%d", B, A, M);
    K7_REQUIRE(generation >= 1u && generation <= K7_BEST_GEN_MAX, "k7_match_scatter: generation %u outside [1, %u]",
               generation, K7_BEST_GEN_MAX);
    K7_REQUIRE(params->neg_thr > 1e-3f && params->neg_thr < 1.0f && params->pos_thr >= params->neg_thr,
               "k7_match_scatter: needs 1e-3 < neg_thr <= pos_thr, neg_thr < 1 (got %f, %f): use k7_match",
               (double)params->neg_thr, (double)params->pos_thr);
    K7_REQUIRE((reinterpret_cast<uintptr_t>(best) & 7) == 0, "k7_match_scatter: best must be 8-byte aligned");
    int rc = k7_check_spec(spec, A);
    if (rc) return rc;
    K7ScatterAux X;
    memset(&X, 0, sizeof(X));
    static_assert(sizeof(K7ScatterShape) == 80, "K7ScatterShape is copied as uint4s");
    K7ScatterShape* shapes = X.shapes;
    const double t = (double)params->neg_thr;
    X.c_in = (float)(t / (1.0 + t) * (1.0 - 1e-4));
    const bool xy = spec->box_order == K7_BOX_XYXY;
    const int K = spec->cell_anchors;
    for (int l = 0; l < spec->num_levels; ++l) {
        X.inv_stride[l] = (float)(1.0 / spec->stride[l]);
        X.lvl_lo[l] = 3.0e38f;
        X.lvl_hi[l] = 0.0f;
        for (int k = 0; k < K; ++k) {
            const double* b = spec->base[l][k];
            K7ScatterShape& S = shapes[l * K + k];
            const double x1 = xy ? b[0] : b[1], y1 = xy ? b[1] : b[0], x2 = xy ? b[2] : b[3], y2 = xy ? b[3] : b[2];
            const double w = x2 - x1, h = y2 - y1;
            for (int c = 0; c < 4; ++c) S.base[c] = b[c];
            // outward roundings: the device compares against float32-rounded anchors (coordinates up to ~1e4: 1e-3 px)
            S.wa = (float)(w * (1.0 + 1e-5) + 2e-3);
            S.ha = (float)(h * (1.0 + 1e-5) + 2e-3);
            const double ar_lo = (w - 2e-3 > 0 ? w - 2e-3 : 0) * (h - 2e-3 > 0 ? h - 2e-3 : 0) * (1.0 - 1e-5);
            const double ar_hi = (w + 2e-3) * (h + 2e-3) * (1.0 + 1e-5);
            S.area_lo = (float)ar_lo;
            S.g_lo = (float)(t * (1.0 - 1e-4) * ar_lo);
            S.g_hi = (float)(ar_hi / (t * (1.0 - 1e-4)));
            S.bx1 = (float)x1; S.bx2 = (float)x2; S.by1 = (float)y1; S.by2 = (float)y2;
            if (S.g_lo < X.lvl_lo[l]) X.lvl_lo[l] = S.g_lo;
            if (S.g_hi > X.lvl_hi[l]) X.lvl_hi[l] = S.g_hi;
        }
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t tbytes = sizeof(K7ScatterShape) * (size_t)spec->num_levels * K + sizeof(K7ScatterLevel) * K7_MAX_LEVELS;
    const long long tasks = (long long)B * M;
    const int wpb = K7_SCATTER_THREADS / 32;
    long long blocks = (tasks + wpb - 1) / wpb;
    const long long cap = 12ll * k7_sm_count();      // 12 resident CTAs of 128 threads per SM; the rest is looped
    if (blocks > cap) blocks = cap;
    if (zero_fill4 && blocks < 4ll * k7_sm_count()) blocks = 4ll * k7_sm_count();   // enough threads for the zero fill
    unsigned char* raw = reinterpret_cast<unsigned char*>(best);
    unsigned int* tcount = reinterpret_cast<unsigned int*>(raw + k7_best_list_off(B, A));
    unsigned int* tlist = reinterpret_cast<unsigned int*>(raw + k7_best_list_off(B, A) + k7_best_list_hdr(B));
    K7_CUDA(cudaMemsetAsync(tcount, 0, k7_best_list_hdr(B), st));
    k7_match_scatter_kernel<<<(unsigned)blocks, K7_SCATTER_THREADS, tbytes, st>>>(
        *spec, X, annotations, B, A, M, params->neg_thr, reinterpret_cast<unsigned long long*>(best), generation,
        tcount, tlist, reinterpret_cast<float4*>(zero_fill4));
    K7_LAUNCH_CHECK();
    return K7_OK;
}

// ------------------------------------------------------------------ ragged annotations -> [B,M,5]
// datasettool/dataloader.py:678-686 pads every image's annotation rows to the batch maximum with -1
// in a Python loop on the host.  Here the batch travels as one ragged block (rows [N,5] + offsets
// [B+1]) and the padded tensor the hot path expects is written on the device.
__global__ void __launch_bounds__(256)
k7_annotations_pad_kernel(const float* __restrict__ rows, const int32_t* __restrict__ offsets, int M,
                          float* __restrict__ out) {
    const int b = blockIdx.y;
    const int lo = offsets[b], n = offsets[b + 1] - lo;
    const int e = blockIdx.x * 256 + threadIdx.x;   // element of the image's [M,5] block
    if (e >= M * 5) return;
    const int m = e / 5;
    out[(size_t)b * M * 5 + e] = (m < n) ? __ldg(rows + (size_t)lo * 5 + e) : -1.0f;
}

extern "C" int k7_annotations_pad(const float* rows, const int32_t* offsets, int B, int M, float* out, void* stream) {
    K7_REQUIRE(offsets && out, "k7_annotations_pad: null argument");
    K7_REQUIRE(B > 0 && M > 0 && B <= 65535 && M < (1 << 27), "k7_annotations_pad: bad shape B=%d M=%d", B, M);
    dim3 grid((M * 5 + 255) / 256, B);
    k7_annotations_pad_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(rows, offsets, M, out);
    K7_LAUNCH_CHECK();
    return K7_OK;
}
