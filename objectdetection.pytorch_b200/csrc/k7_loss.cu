// Focal + smooth-L1 loss (and its gradient) in one streaming pass (training side, stage 2).
//
//   k7_loss   FocalLoss.forward after assignment   architectures/losses.py:62, 94-168
//             (+ loss_efficientdet.py:61-160 through K7LossParams) and its autograd backward.
//
// The job is "move A*C*4 bytes per image in (and, with gradients, out) at HBM speed".  Profiling
// history (profiles/): v1 (row-mapped lanes, per-warp rings) was issue-bound at 58 issue slots per
// element; v2 (flat 4 KB tiles per warp, packed math) reached 27, of which 11 were per-tile
// bookkeeping; this version (v3) moves the bookkeeping out of the compute warps:
//
//   * WARP-SPECIALISED PIPELINE.  One CTA per SM.  A producer warp cuts classifications[b] (a flat
//     array of A*C floats; row boundaries are ignored) into tiles of 16 x 512 floats and drives a
//     ring of shared-memory stages with 1-D bulk async copies (TMA, SASS UBLKCP): tile in, the
//     assignment words of the rows touching it in, and - once the 16 consumer warps have signalled
//     the stage done - gradients out.  full[]/done[] mbarriers; no __syncthreads in the loop.
//   * FLAT PASS.  Consumer warp w owns floats [512w, 512w+512) of every tile: LDS.128 -> math ->
//     STS.128 on float4 number lane+32k (conflict free, perfectly balanced whatever C is).  All
//     elements are evaluated as "target 0" (negative / non-label class) with packed f32x2
//     arithmetic (FFMA2/FMUL2, new on sm_100):
//         p = clamp(x); q = fl(1-p); pp = 1-q; nl = pp*S(pp) ~ -log(q)      (degree-5 minimax)
//         loss += p*p*nl                                  (x (1-alpha) once per image)
//         grad  = (p == x) ? (1-alpha)*scale * p * (2*nl + p/q) : 0          (MUFU.RCP)
//     about 13 issue slots per element.  The polynomial is valid for p <= 0.25; the pass tracks
//     max(p) and a slice that saw a larger value (rare: confident predictions) is redone from
//     global memory by an exact scalar path with logf.
//   * ROW EXCEPTIONS.  Ignored rows (state IGNORE, ~1-2 %) are turned into a bit mask over the
//     slice before the flat pass (only slices that have one take the masked variant); the single
//     "target 1" element of each positive row is patched after the flat pass (its "target 0"
//     contribution is removed and the focal-positive term / gradient written); smooth-L1 and
//     d(loss)/d(reg) are computed by the lane owning a positive row, in the slice where the row
//     starts.  d(reg) of non-positive rows is zero: pre-zeroed by k7_match or by a memset here.
//   * Loss sums stay in registers while a warp works on one image; at an image change they are
//     reduced with shuffles into one float2 per (image, CTA, warp) slot - no atomics - and summed
//     per image in float64 in slot order by the finalize kernel: bit-reproducible run to run.
//
// Tiles never straddle images.  When A*C is not a multiple of 4 the image bases are not 16-byte
// aligned and bulk copies are impossible: a compatibility kernel moves 1024-float tiles with plain
// loads/stores, one warp per tile (odd test shapes only; every real head has A*C % 4 == 0).
//
// Algorithmic bytes per image: forward A*(4C+16)+20M, forward+backward A*(8C+32)+20M.
#include <stdlib.h>

#include "k7_common.cuh"

#define K7_LOSS_MAX_WARPS 16
#define K7_LOSS_TILE 1024   // floats per tile

struct K7LossDev {
    float alpha, one_m_alpha;
    float beta, quad, shift;
    float div[4];
    int use_div;
    int yx_targets;
    int B, A, C, M;
};

// per-image gradient scales, from the match stage's num_pos:
//   x: 1 / (max(num_pos,1) * mean_divisor)      (losses.py:118,168)
//   y: 1 / (4 * num_pos * mean_divisor) or 0    (losses.py:161,168)
__global__ void k7_loss_scales_kernel(const int32_t* __restrict__ num_pos, int B, float mean_div,
                                      float2* __restrict__ scales) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const int n = num_pos[b];
    float2 s;
    s.x = 1.0f / (fmaxf((float)n, 1.0f) * mean_div);
    s.y = n > 0 ? 1.0f / (4.0f * (float)n * mean_div) : 0.0f;
    scales[b] = s;
}

// ------------------------------------------------------------------ element math
__device__ __forceinline__ float k7_rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// Two "target 0" elements.  Returns their gradients (GRAD), accumulates p^2 * -log(1-p) into acc
// and max(p) into pmax.  ign: bit 0 / bit 1 = element belongs to an ignored row.
template <bool GRAD, bool MASKED>
__device__ __forceinline__ float2 k7_neg_pair(const float2 x, const uint32_t ign, const float2 sco, float2& acc,
                                              float& pmax) {
    float2 p;
    p.x = fminf(fmaxf(x.x, K7_P_LO), K7_P_HI);
    p.y = fminf(fmaxf(x.y, K7_P_LO), K7_P_HI);
    pmax = fmaxf(pmax, fmaxf(p.x, p.y));
    const float2 one = make_float2(1.0f, 1.0f), neg1 = make_float2(-1.0f, -1.0f);
    const float2 q = __ffma2_rn(p, neg1, one);     // fl(1 - p), what the reference feeds to log (losses.py:110)
    const float2 pp = __ffma2_rn(q, neg1, one);    // 1 - q, exact
    float2 s = make_float2(0.3357100188732147f, 0.3357100188732147f);
    s = __ffma2_rn(s, pp, make_float2(0.1467030793428421f, 0.1467030793428421f));
    s = __ffma2_rn(s, pp, make_float2(0.25711026787757874f, 0.25711026787757874f));
    s = __ffma2_rn(s, pp, make_float2(0.33293071389198303f, 0.33293071389198303f));
    s = __ffma2_rn(s, pp, make_float2(0.5000072717666626f, 0.5000072717666626f));
    s = __ffma2_rn(s, pp, one);
    const float2 nl = __fmul2_rn(pp, s);           // -log(q)
    float2 t = __fmul2_rn(p, p);
    if (MASKED) {
        if (ign & 1u) t.x = 0.0f;
        if (ign & 2u) t.y = 0.0f;
    }
    acc = __ffma2_rn(t, nl, acc);
    float2 g = make_float2(0.0f, 0.0f);
    if (GRAD) {
        float2 r;
        r.x = k7_rcp_approx(q.x);
        r.y = k7_rcp_approx(q.y);
        const float2 u = __fmul2_rn(p, r);                                   // p / (1-p)
        const float2 v = __ffma2_rn(make_float2(2.0f, 2.0f), nl, u);         // 2*nl + p/(1-p)
        const float2 w = __fmul2_rn(p, sco);                                 // (1-alpha) * scale * p
        g = __fmul2_rn(w, v);
        // torch.clamp passes the gradient on the closed interval  <=>  clamp(x) == x
        const bool k0 = (p.x == x.x) && !(MASKED && (ign & 1u));
        const bool k1 = (p.y == x.y) && !(MASKED && (ign & 2u));
        g.x = k0 ? g.x : 0.0f;
        g.y = k1 ? g.y : 0.0f;
    }
    return g;
}

template <bool GRAD, bool MASKED>
__device__ __forceinline__ void k7_flat4(float4* __restrict__ p4, const uint32_t nib, const float2 sco, float2& acc,
                                         float& pmax) {
    const float4 x = *p4;
    const float2 ga = k7_neg_pair<GRAD, MASKED>(make_float2(x.x, x.y), nib & 3u, sco, acc, pmax);
    const float2 gb = k7_neg_pair<GRAD, MASKED>(make_float2(x.z, x.w), (nib >> 2) & 3u, sco, acc, pmax);
    if (GRAD) *p4 = make_float4(ga.x, ga.y, gb.x, gb.y);
}

// The flat pass over n4 float4s of a tile.  mask: one bit per element (set = ignored row).
template <int TILE_F, bool GRAD, bool MASKED>
__device__ __forceinline__ void k7_flat_pass(float* __restrict__ tile, const int n4, const uint32_t* __restrict__ mask,
                                             const int lane, const float sco1, float2& acc, float& pmax) {
    float4* t4 = reinterpret_cast<float4*>(tile);
    const float2 sco = make_float2(sco1, sco1);
    const int sh = (lane & 7) * 4;
    if (n4 == TILE_F / 4) {
#pragma unroll
        for (int k = 0; k < TILE_F / 128; ++k) {
            uint32_t nib = 0;
            if (MASKED) nib = (mask[(lane >> 3) + 4 * k] >> sh) & 15u;
            k7_flat4<GRAD, MASKED>(t4 + lane + 32 * k, nib, sco, acc, pmax);
        }
    } else {
        for (int f = lane; f < n4; f += 32) {
            uint32_t nib = 0;
            if (MASKED) nib = (mask[f >> 3] >> sh) & 15u;
            k7_flat4<GRAD, MASKED>(t4 + f, nib, sco, acc, pmax);
        }
    }
}

// Exact scalar evaluation of one "target 0" element (any p): loss without the (1-alpha) factor,
// gradient without (1-alpha)*scale.  Used by the redo path and by the label patch.
__device__ __forceinline__ float k7_neg_exact(const float x, float* grad) {
    const float p = k7_clamp_p(x);
    const float q = 1.0f - p;
    const float nl = k7_neg_log1m(q);
    if (grad) *grad = (p == x) ? p * fmaf(2.0f, nl, p * k7_rcp_approx(q)) : 0.0f;
    return (p * p) * nl;
}

// ------------------------------------------------------------------ row helpers (shared by both kernels)
// Smooth-L1 of positive row `a` of image b (losses.py:122-166): returns the loss, writes d(reg).
template <bool GRAD, bool FROM_SPEC>
__device__ __forceinline__ float k7_row_smooth_l1(const K7AnchorSpec& spec, const float* __restrict__ anchors,
                                                  int box_order, const K7LossDev& P, int a, size_t R,
                                                  const float* __restrict__ arow, const float* __restrict__ reg,
                                                  float scale_y, float4* __restrict__ dreg) {
    float ax1, ay1, ax2, ay2;
    k7_anchor<FROM_SPEC>(spec, anchors, box_order, a, ax1, ay1, ax2, ay2);
    const float g0 = __ldg(arow + 0), g1 = __ldg(arow + 1), g2 = __ldg(arow + 2), g3 = __ldg(arow + 3);
    const K7Reg4 tg = k7_reg_targets(ax1, ay1, ax2, ay2, g0, g1, g2, g3, P.use_div ? P.div : nullptr, P.yx_targets);
    const float4 rv = __ldg(reinterpret_cast<const float4*>(reg) + R);
    float gx, gy, gz, gw;
    float l = k7_smooth_l1(tg.v[0], rv.x, P.beta, P.quad, P.shift, &gx);
    l += k7_smooth_l1(tg.v[1], rv.y, P.beta, P.quad, P.shift, &gy);
    l += k7_smooth_l1(tg.v[2], rv.z, P.beta, P.quad, P.shift, &gz);
    l += k7_smooth_l1(tg.v[3], rv.w, P.beta, P.quad, P.shift, &gw);
    if (GRAD) dreg[R] = make_float4(gx * scale_y, gy * scale_y, gz * scale_y, gw * scale_y);
    return l;
}

// bits [lo, hi) of a span of elements -> the 32-bit mask word whose first element is `base`
__device__ __forceinline__ uint32_t k7_range_bits(int lo, int hi, int base) {
    const int a = max(lo - base, 0), z = min(hi - base, 32);
    if (z <= a) return 0u;
    return (z - a == 32) ? 0xffffffffu : (((1u << (z - a)) - 1u) << a);
}

// ------------------------------------------------------------------ the pipelined kernel (A*C % 4 == 0)
// One CTA per SM.  Warp CW is the PRODUCER: one elected lane cuts the image into tiles of
// CW*512 floats, issues the TMA loads (classification tile + the assignment words of the rows it
// touches) into a ring of NS stages, and, once the CW consumer warps have signalled a stage
// done, the TMA store of the gradients.  Warps 0..CW-1 are CONSUMERS: warp w owns floats
// [512w, 512w+512) of every tile.  full[s] / done[s] are mbarriers; there is no __syncthreads
// after the prologue.
#define K7_PIPE_SLICE 512
#define K7_PIPE_MAX_CW 16

struct K7PipeGeom {
    uint32_t AC;          // elements per image
    uint32_t tile_f;      // floats per tile = cw * K7_PIPE_SLICE
    uint32_t tpi;         // tiles per image
    uint32_t num_tiles;   // B * tpi
    K7FastDiv div_c, div_tpi;
    uint32_t side_words;  // assignment words reserved per stage
    uint32_t stage_stride;
    uint32_t off_desc, off_mask, off_bars;
    int stages, cw, grid;
    size_t smem_bytes;
};

struct K7TileDesc {   // written by the producer next to each stage, 32 bytes
    uint32_t n, rem, lead, b;
    uint32_t row_first, off, pad0, pad1;
};

template <bool GRAD, bool FROM_SPEC>
__global__ void __launch_bounds__((K7_PIPE_MAX_CW + 1) * 32, 1)
k7_loss_pipe_kernel(const __grid_constant__ K7AnchorSpec spec, const float* __restrict__ anchors, int box_order,
                    const float* __restrict__ cls, const float* __restrict__ reg, const float* __restrict__ annotations,
                    const int32_t* __restrict__ assign, const float2* __restrict__ scales,
                    const __grid_constant__ K7LossDev P, const __grid_constant__ K7PipeGeom G,
                    float2* __restrict__ partials, float* __restrict__ dcls, float4* __restrict__ dreg) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int NS = G.stages, CW = G.cw;
    const int C = P.C, A = P.A;
    constexpr int SL = K7_PIPE_SLICE;
    K7TileDesc* desc = reinterpret_cast<K7TileDesc*>(smem + G.off_desc);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + G.off_bars);
    uint64_t* done = full + NS;

    if (threadIdx.x == 0) {
        for (int s = 0; s < NS; ++s) {
            k7_mbar_init(full + s, 1);
            k7_mbar_init(done + s, (uint32_t)CW);
        }
        k7_fence_mbar_init();
        k7_fence_proxy_async();
    }
    __syncthreads();
    if (blockIdx.x >= G.num_tiles) return;
    const uint32_t ntl = (G.num_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x;   // tiles of this CTA

    if (warp == CW) {
        // ================================================================ producer
        if (lane != 0) return;
        auto load = [&](const uint32_t j) {
            const uint32_t g = blockIdx.x + j * gridDim.x;
            const int s = (int)(j % (uint32_t)NS);
            unsigned char* stage = smem + (size_t)s * G.stage_stride;
            K7TileDesc d;
            d.b = k7_div(g, G.div_tpi);
            d.off = (g - d.b * G.tpi) * G.tile_f;
            d.n = min(G.tile_f, G.AC - d.off);
            d.row_first = k7_div(d.off, G.div_c);
            d.rem = d.off - d.row_first * (uint32_t)C;
            const uint32_t nrows = k7_div(d.rem + d.n - 1u, G.div_c) + 1u;
            const size_t g0 = (size_t)d.b * A + d.row_first;
            d.lead = (uint32_t)(g0 & 3);
            d.pad0 = d.pad1 = 0;
            const uint32_t side_bytes = ((d.lead + nrows + 3u) & ~3u) * 4u;
            desc[s] = d;
            k7_mbar_expect_tx(full + s, d.n * 4u + side_bytes);   // release: orders the descriptor
            k7_bulk_g2s(stage, cls + (size_t)d.b * G.AC + d.off, d.n * 4u, full + s);
            k7_bulk_g2s(stage + (size_t)G.tile_f * 4, assign + (g0 - d.lead), side_bytes, full + s);
        };
        // with gradients one stage is always draining its store
        const uint32_t ahead = (uint32_t)(GRAD ? NS - 1 : NS);
        uint32_t jl = 0;
        for (; jl < ntl && jl < ahead; ++jl) load(jl);
        int s = 0;
        uint32_t phase = 0;
        for (uint32_t j = 0; j < ntl; ++j) {
            k7_mbar_wait(done + s, phase);
            if (GRAD) {
                const K7TileDesc d = desc[s];
                k7_bulk_s2g(dcls + (size_t)d.b * G.AC + d.off, smem + (size_t)s * G.stage_stride, d.n * 4u);
                k7_bulk_commit();
                if (jl < ntl) {
                    k7_bulk_wait_read<1>();   // the store of tile j-1 has left its stage
                    load(jl++);
                }
            } else if (jl < ntl) {
                load(jl++);
            }
            if (++s == NS) {
                s = 0;
                phase ^= 1u;
            }
        }
        if (GRAD) k7_bulk_wait<0>();
        return;
    }
    if (warp > CW) return;

    // ==================================================================== consumers
    uint32_t* mask = reinterpret_cast<uint32_t*>(smem + G.off_mask) + warp * (SL / 32);
    float2* my_partials = partials + (size_t)blockIdx.x * CW + warp;      // + b * (gridDim.x * CW)
    const size_t pstride = (size_t)gridDim.x * CW;
    for (int b = lane; b < P.B; b += 32) my_partials[(size_t)b * pstride] = make_float2(0.f, 0.f);
    __syncwarp();

    const float oma = P.one_m_alpha;
    float2 acc = make_float2(0.0f, 0.0f);   // sum p^2 * -log(1-p) of the current image (this lane)
    float cfix = 0.0f, cslow = 0.0f, rsum = 0.0f;
    int cur_b = -1;
    float2 sc = make_float2(0.f, 0.f);
    float sco = 0.0f;
    auto flush = [&]() {
        float c = fmaf(oma, (acc.x + acc.y) + cslow, cfix);
        c = k7_warp_sum(c);
        const float r = k7_warp_sum(rsum);
        if (lane == 0) my_partials[(size_t)cur_b * pstride] = make_float2(c, r);
        acc = make_float2(0.0f, 0.0f);
        cfix = cslow = rsum = 0.0f;
    };
    const int i_lo = warp * SL;

    const uint32_t full_u32 = k7_smem_u32(full), done_u32 = k7_smem_u32(done);
    int s = -1;
    uint32_t phase = 1u;
    for (uint32_t j = 0; j < ntl; ++j) {
        if (++s == NS) s = 0;
        if (s == 0) phase ^= 1u;
        unsigned char* stage = smem + (size_t)s * G.stage_stride;
        k7_mbar_wait_u32(full_u32 + 8u * s, phase);
        const uint4 d0 = *reinterpret_cast<const uint4*>(desc + s);   // n, rem, lead, b
        const int n = (int)d0.x;
        if ((int)d0.w != cur_b) {
            if (cur_b >= 0) flush();
            cur_b = (int)d0.w;
            sc = __ldg(scales + cur_b);
            sco = oma * sc.x;
        }
        if (i_lo < n) {
            const int rem = (int)d0.y;
            const int i_hi = min(n, i_lo + SL);
            const int len = i_hi - i_lo, n4 = len >> 2;
            float* tile = reinterpret_cast<float*>(stage);
            float* slice = tile + i_lo;
            const int32_t* side = reinterpret_cast<const int32_t*>(stage + (size_t)G.tile_f * 4) + d0.z;
            const int k_lo = (int)k7_div((uint32_t)(rem + i_lo), G.div_c);
            const int k_hi = (int)k7_div((uint32_t)(rem + i_hi - 1), G.div_c);

            // ---- scan the rows touching the slice: ignored rows -> element mask, any positive row?
            uint32_t mword = 0u;
            bool masked = false, any_pos = false;
            for (int k0 = k_lo; k0 <= k_hi; k0 += 32) {
                const int k = k0 + lane;
                const int st = (k <= k_hi) ? K7_ASSIGN_STATE(side[k]) : K7_STATE_NEG;
                unsigned ib = __ballot_sync(0xffffffffu, st == K7_STATE_IGNORE);
                any_pos |= __any_sync(0xffffffffu, st == K7_STATE_POS);
                if (ib == 0u) continue;
                masked = true;
                if (C == 1) {   // row k is element k: the ballot IS the mask word
                    if (lane == ((k0 - k_lo) >> 5)) mword = ib;
                } else {
                    while (ib) {
                        const int kk = k0 + __ffs(ib) - 1;
                        ib &= ib - 1;
                        const int lo = kk * C - rem - i_lo;
                        mword |= k7_range_bits(max(lo, 0), min(lo + C, len), 32 * lane);
                    }
                }
            }
            if (masked) {
                if (lane < SL / 32) mask[lane] = mword;
                __syncwarp();
            }

            // ---- flat pass
            const float2 acc0 = acc;
            float pmax = 0.0f;
            if (masked)
                k7_flat_pass<SL, GRAD, true>(slice, n4, mask, lane, sco, acc, pmax);
            else
                k7_flat_pass<SL, GRAD, false>(slice, n4, mask, lane, sco, acc, pmax);
            if (__any_sync(0xffffffffu, pmax > K7_POLY_MAX)) {
                // a confident prediction: outside the polynomial's range.  Redo the slice exactly.
                __syncwarp();
                acc = acc0;
                const float* src = cls + (size_t)d0.w * G.AC + desc[s].off + i_lo;
                for (int i = lane; i < len; i += 32) {
                    const bool ign = masked && ((mask[i >> 5] >> (i & 31)) & 1u);
                    float gr;
                    const float l = k7_neg_exact(__ldg(src + i), &gr);
                    if (!ign) cslow += l;
                    if (GRAD) slice[i] = ign ? 0.0f : sco * gr;
                }
            }

            // ---- positive rows: smooth-L1 / d(reg) where the row starts, label patch where the label lies
            if (any_pos) {
                __syncwarp();
                const uint32_t row_first = desc[s].row_first, off = desc[s].off;
                const size_t e0 = (size_t)d0.w * G.AC + off;
                for (int k = k_lo + lane; k <= k_hi; k += 32) {
                    const int aw = side[k];
                    if (K7_ASSIGN_STATE(aw) != K7_STATE_POS) continue;
                    const float* arow = annotations + ((size_t)cur_b * P.M + K7_ASSIGN_ROW(aw)) * 5;
                    const int i0 = k * C - rem;   // tile-local element index of the row's class 0
                    if (i0 >= i_lo && i0 < i_hi) {
                        const int a = (int)row_first + k;
                        rsum += k7_row_smooth_l1<GRAD, FROM_SPEC>(spec, anchors, box_order, P, a, (size_t)cur_b * A + a,
                                                                  arow, reg, sc.y, dreg);
                    }
                    const int label = (int)__ldg(arow + 4);   // .long() truncation, losses.py:97
                    const int il = i0 + label;
                    if (label >= 0 && label < C && il >= i_lo && il < i_hi) {
                        // the label class has target 1: replace its "target 0" term
                        const float xl = __ldg(cls + e0 + il);
                        const float p = k7_clamp_p(xl);
                        cfix += k7_focal_pos(p, P.alpha) - oma * k7_neg_exact(xl, nullptr);
                        if (GRAD) tile[il] = (p == xl) ? k7_focal_pos_grad(p, P.alpha) * sc.x : 0.0f;
                    }
                }
            }
        }
        if (GRAD) k7_fence_proxy_async();
        __syncwarp();
        if (lane == 0) k7_mbar_arrive_u32(done_u32 + 8u * s);
    }
    if (cur_b >= 0) flush();
}

// ------------------------------------------------------------------ the compatibility kernel (A*C % 4 != 0)
// Image bases are not 16-byte aligned, so no bulk copies: one warp per 1024-float tile moves it
// with plain loads/stores.  Same flat pass, same row handling, per-tile partials.
struct K7Tile {
    uint32_t b, off, n, row_first, rem, nrows;
};
struct K7CompatGeom {
    uint32_t AC, tpi, num_tiles;
    K7FastDiv div_c, div_tpi;
    uint32_t stage_stride;
    int warps, grid;
    size_t smem_bytes;
};

template <int TILE_F, bool GRAD, bool FROM_SPEC>
__global__ void __launch_bounds__(K7_LOSS_MAX_WARPS * 32, 1)
k7_loss_compat_kernel(const __grid_constant__ K7AnchorSpec spec, const float* __restrict__ anchors, int box_order,
                      const float* __restrict__ cls, const float* __restrict__ reg, const float* __restrict__ annotations,
                      const int32_t* __restrict__ assign, const float2* __restrict__ scales,
                      const __grid_constant__ K7LossDev P, const __grid_constant__ K7CompatGeom G,
                      float2* __restrict__ partials, float* __restrict__ dcls, float4* __restrict__ dreg) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int C = P.C, A = P.A;
    unsigned char* stage = smem + (size_t)warp * G.stage_stride;
    float* tile = reinterpret_cast<float*>(stage);
    int32_t* side = reinterpret_cast<int32_t*>(stage + TILE_F * 4);
    uint32_t* mask = reinterpret_cast<uint32_t*>(smem + (size_t)G.warps * G.stage_stride) + warp * (TILE_F / 32);
    const uint32_t nW = gridDim.x * G.warps;
    const float oma = P.one_m_alpha;

    for (uint64_t g64 = blockIdx.x * G.warps + warp; g64 < G.num_tiles; g64 += nW) {
        const uint32_t g = (uint32_t)g64;
        K7Tile t;
        t.b = k7_div(g, G.div_tpi);
        t.off = (g - t.b * G.tpi) * (uint32_t)TILE_F;
        t.n = min((uint32_t)TILE_F, G.AC - t.off);
        t.row_first = k7_div(t.off, G.div_c);
        t.rem = t.off - t.row_first * (uint32_t)C;
        t.nrows = k7_div(t.rem + t.n - 1u, G.div_c) + 1u;
        const size_t e0 = (size_t)t.b * G.AC + t.off;     // first element of the tile in cls / dcls
        const size_t g0 = (size_t)t.b * A + t.row_first;  // first row of the tile in assign / reg / dreg
        const int n = (int)t.n, nrows = (int)t.nrows, rem = (int)t.rem;
        const int n4 = (n + 3) >> 2;
        for (int i = lane; i < 4 * n4; i += 32) tile[i] = (i < n) ? cls[e0 + i] : 0.0f;
        for (int k = lane; k < nrows; k += 32) side[k] = assign[g0 + k];
        __syncwarp();

        uint32_t mword = 0u;
        bool masked = false;
        for (int k0 = 0; k0 < nrows; k0 += 32) {
            const int k = k0 + lane;
            const int st = (k < nrows) ? K7_ASSIGN_STATE(side[k]) : K7_STATE_NEG;
            unsigned ib = __ballot_sync(0xffffffffu, st == K7_STATE_IGNORE);
            if (ib == 0u) continue;
            masked = true;
            while (ib) {
                const int kk = k0 + __ffs(ib) - 1;
                ib &= ib - 1;
                const int lo = kk * C - rem;
                mword |= k7_range_bits(max(lo, 0), min(lo + C, n), 32 * lane);
            }
        }
        if (n & 3) {   // zero padding of the last float4
            masked = true;
            mword |= k7_range_bits(n, 4 * n4, 32 * lane);
        }
        if (masked) {
            mask[lane] = mword;
            __syncwarp();
        }

        const float2 sc = __ldg(scales + t.b);
        const float sco = oma * sc.x;
        float2 acc = make_float2(0.0f, 0.0f);
        float pmax = 0.0f;
        if (masked)
            k7_flat_pass<TILE_F, GRAD, true>(tile, n4, mask, lane, sco, acc, pmax);
        else
            k7_flat_pass<TILE_F, GRAD, false>(tile, n4, mask, lane, sco, acc, pmax);
        float csum = acc.x + acc.y;
        if (__any_sync(0xffffffffu, pmax > K7_POLY_MAX)) {
            __syncwarp();
            csum = 0.0f;
            for (int i = lane; i < n; i += 32) {
                const bool ign = masked && ((mask[i >> 5] >> (i & 31)) & 1u);
                float gr;
                const float l = k7_neg_exact(__ldg(cls + e0 + i), &gr);
                if (!ign) csum += l;
                if (GRAD) tile[i] = ign ? 0.0f : sco * gr;
            }
        }
        csum *= oma;
        __syncwarp();

        float rsum = 0.0f;
        for (int k = lane; k < nrows; k += 32) {
            const int aw = side[k];
            const int i0 = k * C - rem;
            const bool starts = i0 >= 0;
            const size_t R = g0 + k;
            bool wrote = false;
            if (K7_ASSIGN_STATE(aw) == K7_STATE_POS) {
                const float* arow = annotations + ((size_t)t.b * P.M + K7_ASSIGN_ROW(aw)) * 5;
                if (starts) {
                    rsum += k7_row_smooth_l1<GRAD, FROM_SPEC>(spec, anchors, box_order, P, (int)(t.row_first + k), R, arow,
                                                              reg, sc.y, dreg);
                    wrote = true;
                }
                const int label = (int)__ldg(arow + 4);
                const int il = i0 + label;
                if (label >= 0 && label < C && il >= 0 && il < n) {
                    const float xl = __ldg(cls + e0 + il);
                    const float p = k7_clamp_p(xl);
                    csum += k7_focal_pos(p, P.alpha) - oma * k7_neg_exact(xl, nullptr);
                    if (GRAD) tile[il] = (p == xl) ? k7_focal_pos_grad(p, P.alpha) * sc.x : 0.0f;
                }
            }
            if (GRAD && starts && !wrote) dreg[R] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        csum = k7_warp_sum(csum);
        rsum = k7_warp_sum(rsum);
        if (lane == 0) partials[g] = make_float2(csum, rsum);
        __syncwarp();
        if (GRAD) {
            for (int i = lane; i < n; i += 32) dcls[e0 + i] = tile[i];
            __syncwarp();
        }
    }
}

// Per image: sum the tile partials in float64, fixed order; write (cls_loss_i, reg_loss_i).
// The last block to finish reduces the batch (threadfence + counter).
__global__ void __launch_bounds__(256)
k7_loss_finalize_kernel(const float2* __restrict__ partials, const int32_t* __restrict__ num_pos, int B, uint32_t tpi,
                        float* __restrict__ per_image, float* __restrict__ sums, unsigned int* __restrict__ counter) {
    __shared__ double s_c[256], s_r[256], s_n[256];
    __shared__ bool s_last;
    const int b = blockIdx.x;
    const float2* pt = partials + (size_t)b * tpi;
    double c = 0.0, r = 0.0;
    for (uint32_t t = threadIdx.x; t < tpi; t += blockDim.x) {
        const float2 p = pt[t];
        c += p.x;
        r += p.y;
    }
    s_c[threadIdx.x] = c;
    s_r[threadIdx.x] = r;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) {
            s_c[threadIdx.x] += s_c[threadIdx.x + o];
            s_r[threadIdx.x] += s_r[threadIdx.x + o];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const int n = num_pos[b];
        const float ci = (float)s_c[0] / fmaxf((float)n, 1.0f);
        const float ri = n > 0 ? (float)(s_r[0] / (4.0 * (double)n)) : 0.0f;
        per_image[2 * b + 0] = ci;
        per_image[2 * b + 1] = ri;
        __threadfence();
        const unsigned int done = atomicAdd(counter, 1u);
        s_last = (done == (unsigned int)(B - 1));
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        // batch sums in image order (deterministic); 256 threads, tree in shared memory
        double cc = 0.0, rr = 0.0, nn = 0.0;
        for (int i = threadIdx.x; i < B; i += blockDim.x) {
            cc += (double)__ldcg(per_image + 2 * i);
            rr += (double)__ldcg(per_image + 2 * i + 1);
            nn += (double)num_pos[i];
        }
        s_c[threadIdx.x] = cc;
        s_r[threadIdx.x] = rr;
        s_n[threadIdx.x] = nn;
        __syncthreads();
        for (int o = 128; o > 0; o >>= 1) {
            if (threadIdx.x < o) {
                s_c[threadIdx.x] += s_c[threadIdx.x + o];
                s_r[threadIdx.x] += s_r[threadIdx.x + o];
                s_n[threadIdx.x] += s_n[threadIdx.x + o];
            }
            __syncthreads();
        }
        if (threadIdx.x == 0) {
            sums[0] = (float)s_c[0];
            sums[1] = (float)s_r[0];
            sums[2] = (float)s_n[0];
            sums[3] = (float)B;
            *counter = 0u;   // re-arm for the next call on this workspace
        }
    }
}

__global__ void k7_scale_kernel(float* __restrict__ g, size_t n, const float* __restrict__ scale) {
    const float s = __ldg(scale);
    if (s == 1.0f) return;
    const size_t n4 = n >> 2;
    float4* g4 = reinterpret_cast<float4*>(g);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 v = g4[i];
        v.x *= s; v.y *= s; v.z *= s; v.w *= s;
        g4[i] = v;
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) g[(n4 << 2) + threadIdx.x] *= s;
}

extern "C" int k7_scale_inplace(float* grad, size_t n, const float* scale, void* stream) {
    K7_REQUIRE(grad && scale, "k7_scale_inplace: null argument");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(grad) & 15) == 0, "k7_scale_inplace: grad must be 16-byte aligned");
    if (n == 0) return K7_OK;
    const int blocks = k7_sm_count() * 8;
    k7_scale_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(grad, n, scale);
    K7_LAUNCH_CHECK();
    return K7_OK;
}

// ------------------------------------------------------------------ host side
#define K7_LOSS_MAX_GRID 256   // upper bound on CTAs assumed by the workspace query

static uint32_t k7_loss_tpi(int A, int C) {
    const unsigned long long ac = (unsigned long long)A * (unsigned long long)C;
    return (uint32_t)((ac + K7_LOSS_TILE - 1) / K7_LOSS_TILE);
}

static size_t k7_loss_ws_layout(int B, unsigned long long slots_per_image, size_t* off_scales, size_t* off_partials,
                                size_t* off_counter) {
    size_t o = 0;
    *off_counter = o;
    o += 256;
    *off_scales = o;
    o += k7_align_up(sizeof(float2) * (size_t)B, 256);
    *off_partials = o;
    o += k7_align_up(sizeof(float2) * (size_t)B * slots_per_image, 256);
    return o;
}

extern "C" size_t k7_loss_workspace_bytes(int B, int A, int C) {
    if (B <= 0 || A <= 0 || C <= 0) return 0;
    size_t a, b, c;
    unsigned long long slots = (unsigned long long)K7_LOSS_MAX_GRID * K7_PIPE_MAX_CW;   // pipelined kernel
    if (((unsigned long long)A * C) % 4ull) slots = k7_loss_tpi(A, C);                  // compatibility kernel
    return k7_loss_ws_layout(B, slots, &a, &b, &c);
}

static int k7_env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    if (!v || !*v) return dflt;
    return atoi(v);
}

static int k7_pipe_geometry(int B, int A, int C, bool grad, K7PipeGeom* g) {
    const unsigned long long ac = (unsigned long long)A * (unsigned long long)C;
    K7_REQUIRE(ac < (1ull << 30), "k7_loss: A*C=%llu per image is too large", ac);
    g->AC = (uint32_t)ac;
    int cw = k7_env_int("K7_WARPS", K7_PIPE_MAX_CW);
    if (cw > K7_PIPE_MAX_CW) cw = K7_PIPE_MAX_CW;
    if (cw < 1) cw = 1;
    g->cw = cw;
    g->tile_f = (uint32_t)cw * K7_PIPE_SLICE;
    g->tpi = (uint32_t)((ac + g->tile_f - 1) / g->tile_f);
    const unsigned long long nt = (unsigned long long)B * g->tpi;
    K7_REQUIRE(nt < (1ull << 31), "k7_loss: %llu tiles exceed the 32-bit tile index", nt);
    g->num_tiles = (uint32_t)nt;
    g->div_c = k7_fastdiv_make((uint32_t)C);
    g->div_tpi = k7_fastdiv_make(g->tpi);
    // rows touching a tile: (tile_f-1)/C + 2 at most, + 3 words of alignment lead, rounded to 16 bytes
    g->side_words = (uint32_t)(((g->tile_f - 1) / (uint32_t)C + 2 + 3 + 3) & ~3u);
    g->stage_stride = (uint32_t)k7_align_up((size_t)g->tile_f * 4 + (size_t)g->side_words * 4, 128);
    const size_t budget = 226 * 1024;
    const size_t fixed = 64 /*alignment slack*/ + (size_t)cw * (K7_PIPE_SLICE / 8);
    int stages = (int)((budget - fixed) / ((size_t)g->stage_stride + sizeof(K7TileDesc) + 16));
    const int want = k7_env_int("K7_STAGES", 8);
    if (stages > want) stages = want;
    if (stages > 16) stages = 16;
    K7_REQUIRE(stages >= (grad ? 3 : 2), "k7_loss: C=%d needs %u bytes of shared memory per stage", C, g->stage_stride);
    g->stages = stages;
    size_t o = (size_t)stages * g->stage_stride;
    g->off_desc = (uint32_t)o;
    o += (size_t)stages * sizeof(K7TileDesc);
    g->off_mask = (uint32_t)o;
    o += (size_t)cw * (K7_PIPE_SLICE / 8);
    o = k7_align_up(o, 16);
    g->off_bars = (uint32_t)o;
    o += (size_t)stages * 16;
    g->smem_bytes = o;
    int grid = k7_sm_count();
    if (grid > K7_LOSS_MAX_GRID) grid = K7_LOSS_MAX_GRID;
    if (nt < (unsigned long long)grid) grid = (int)nt;
    g->grid = grid < 1 ? 1 : grid;
    return K7_OK;
}

static int k7_compat_geometry(int B, int A, int C, K7CompatGeom* g) {
    const unsigned long long ac = (unsigned long long)A * (unsigned long long)C;
    K7_REQUIRE(ac < (1ull << 30), "k7_loss: A*C=%llu per image is too large", ac);
    g->AC = (uint32_t)ac;
    g->tpi = k7_loss_tpi(A, C);
    const unsigned long long nt = (unsigned long long)B * g->tpi;
    K7_REQUIRE(nt < (1ull << 31), "k7_loss: %llu tiles exceed the 32-bit tile index", nt);
    g->num_tiles = (uint32_t)nt;
    g->div_c = k7_fastdiv_make((uint32_t)C);
    g->div_tpi = k7_fastdiv_make(g->tpi);
    const uint32_t side_words = (uint32_t)(((K7_LOSS_TILE - 1) / C + 2 + 3) & ~3);
    g->stage_stride = (uint32_t)k7_align_up((size_t)K7_LOSS_TILE * 4 + (size_t)side_words * 4, 128);
    const size_t per_warp = (size_t)g->stage_stride + K7_LOSS_TILE / 8;
    int warps = (int)((200 * 1024) / per_warp);
    if (warps > K7_LOSS_MAX_WARPS) warps = K7_LOSS_MAX_WARPS;
    K7_REQUIRE(warps >= 1, "k7_loss: C=%d needs %zu bytes of shared memory per warp", C, per_warp);
    g->warps = warps;
    g->smem_bytes = (size_t)warps * per_warp;
    unsigned long long ctas = (nt + warps - 1) / warps;
    int grid = k7_sm_count();
    if (ctas < (unsigned long long)grid) grid = (int)ctas;
    g->grid = grid < 1 ? 1 : grid;
    return K7_OK;
}

extern "C" int k7_loss(const K7AnchorSpec* spec, const float* anchors, int box_order, const float* cls, const float* reg,
                       const float* annotations, const int32_t* assign, const int32_t* num_pos, int B, int A, int C, int M,
                       const K7LossParams* params, float* per_image, float* sums, float* dcls, float* dreg,
                       void* workspace, size_t workspace_bytes, void* stream) {
    K7_REQUIRE(cls && reg && annotations && assign && num_pos && params && per_image && sums && workspace,
               "k7_loss: null argument");
    K7_REQUIRE(spec || anchors, "k7_loss: need an anchor spec or an anchor tensor");
    K7_REQUIRE((dcls == nullptr) == (dreg == nullptr), "k7_loss: dcls and dreg must be given together");
    K7_REQUIRE(B > 0 && A > 0 && C > 0 && M > 0, "k7_loss: bad shape B=%d A=%d C=%d M=%d", B, A, C, M);
    K7_REQUIRE(params->mean_divisor > 0, "k7_loss: mean_divisor must be positive");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(cls) & 15) == 0 && (reinterpret_cast<uintptr_t>(reg) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(assign) & 15) == 0,
               "k7_loss: cls, reg and assign must be 16-byte aligned");
    K7_REQUIRE(!dcls || ((reinterpret_cast<uintptr_t>(dcls) & 15) == 0 && (reinterpret_cast<uintptr_t>(dreg) & 15) == 0),
               "k7_loss: dcls and dreg must be 16-byte aligned");
    if (spec) {
        int rc = k7_check_spec(spec, A);
        if (rc) return rc;
        box_order = spec->box_order;
    } else {
        K7_REQUIRE((reinterpret_cast<uintptr_t>(anchors) & 15) == 0, "k7_loss: anchors must be 16-byte aligned");
        K7_REQUIRE(box_order == K7_BOX_XYXY || box_order == K7_BOX_YXYX, "k7_loss: bad box_order");
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool grad = dcls != nullptr;
    const bool pipe = (((unsigned long long)A * C) % 4ull) == 0ull;

    K7PipeGeom GP;
    K7CompatGeom GC;
    int rc = pipe ? k7_pipe_geometry(B, A, C, grad, &GP) : k7_compat_geometry(B, A, C, &GC);
    if (rc) return rc;
    const unsigned long long slots = pipe ? (unsigned long long)GP.grid * GP.cw : GC.tpi;

    size_t off_scales, off_partials, off_counter;
    const size_t need = k7_loss_ws_layout(B, slots, &off_scales, &off_partials, &off_counter);
    if (workspace_bytes < need) {
        k7_set_error("k7_loss: workspace %zu < %zu bytes", workspace_bytes, need);
        return K7_EWORKSPACE;
    }
    unsigned char* ws = static_cast<unsigned char*>(workspace);
    K7_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "k7_loss: workspace must be 256-byte aligned");
    float2* scales = reinterpret_cast<float2*>(ws + off_scales);
    float2* partials = reinterpret_cast<float2*>(ws + off_partials);
    unsigned int* counter = reinterpret_cast<unsigned int*>(ws + off_counter);

    K7LossDev P;
    P.alpha = params->alpha;
    P.one_m_alpha = 1.0f - params->alpha;   // losses.py:103: 1. - alpha_factor in float32
    P.beta = params->beta;
    P.quad = (float)(0.5 / (double)params->beta);   // 0.5 * 9.0 (losses.py:158)
    P.shift = (float)(0.5 * (double)params->beta);  // 0.5 / 9.0 (losses.py:159)
    if (params->beta == (float)(1.0 / 9.0)) {
        P.quad = 4.5f;
        P.shift = (float)(0.5 / 9.0);
    }
    P.use_div = 0;
    for (int i = 0; i < 4; ++i) {
        P.div[i] = params->target_div[i];
        if (params->target_div[i] != 1.0f) P.use_div = 1;
    }
    P.yx_targets = params->yx_targets;
    P.B = B; P.A = A; P.C = C; P.M = M;

    K7_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned int), st));
    // the pipelined kernel writes d(reg) of positive rows only
    if (grad && pipe && !params->dreg_prezeroed) K7_CUDA(cudaMemsetAsync(dreg, 0, sizeof(float) * 4 * (size_t)B * A, st));
    k7_loss_scales_kernel<<<(B + 127) / 128, 128, 0, st>>>(num_pos, B, (float)params->mean_divisor, scales);
    K7_LAUNCH_CHECK();

    static const K7AnchorSpec kEmpty = {};
    const K7AnchorSpec& sp = spec ? *spec : kEmpty;
    float4* dreg4 = reinterpret_cast<float4*>(dreg);
#define K7_LOSS_ARGS sp, anchors, box_order, cls, reg, annotations, assign, scales, P
    if (pipe) {
#define K7_PIPE_LAUNCH(GR, SP)                                                                                        \
    do {                                                                                                              \
        auto kern = k7_loss_pipe_kernel<GR, SP>;                                                                      \
        K7_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GP.smem_bytes));         \
        kern<<<GP.grid, (GP.cw + 1) * 32, GP.smem_bytes, st>>>(K7_LOSS_ARGS, GP, partials, dcls, dreg4);              \
    } while (0)
        if (grad) { if (spec) K7_PIPE_LAUNCH(true, true); else K7_PIPE_LAUNCH(true, false); }
        else      { if (spec) K7_PIPE_LAUNCH(false, true); else K7_PIPE_LAUNCH(false, false); }
#undef K7_PIPE_LAUNCH
    } else {
#define K7_COMPAT_LAUNCH(GR, SP)                                                                                      \
    do {                                                                                                              \
        auto kern = k7_loss_compat_kernel<K7_LOSS_TILE, GR, SP>;                                                      \
        K7_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GC.smem_bytes));         \
        kern<<<GC.grid, GC.warps * 32, GC.smem_bytes, st>>>(K7_LOSS_ARGS, GC, partials, dcls, dreg4);                 \
    } while (0)
        if (grad) { if (spec) K7_COMPAT_LAUNCH(true, true); else K7_COMPAT_LAUNCH(true, false); }
        else      { if (spec) K7_COMPAT_LAUNCH(false, true); else K7_COMPAT_LAUNCH(false, false); }
#undef K7_COMPAT_LAUNCH
    }
#undef K7_LOSS_ARGS
    K7_LAUNCH_CHECK();

    k7_loss_finalize_kernel<<<B, 256, 0, st>>>(partials, num_pos, B, (uint32_t)slots, per_image, sums, counter);
    K7_LAUNCH_CHECK();
    return K7_OK;
}
