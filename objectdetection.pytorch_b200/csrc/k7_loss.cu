// Focal + smooth-L1 loss (and its gradient) in one streaming pass (training side, stage 2).
//
//   k7_loss   FocalLoss.forward after assignment   architectures/losses.py:62, 94-168
//             (+ loss_efficientdet.py:61-160 through K7LossParams) and its autograd backward.
//
// The job is "move B*A*C*4 bytes in (and, with gradients, out) at HBM speed".  Profiling history
// (profiles/): v1 (row-mapped lanes, per-warp rings) was issue-bound at 58 issue slots per element;
// v2 (flat 4 KB tiles per warp, packed math) reached 27; v3 (warp-specialised CTA pipeline) 13, but
// its consumer warps still served the positive rows inline - dependent global loads that held a
// pipeline stage for ~1 us each; v4 moved the positives into a row kernel behind the stream, fed by an
// exhaustive assignment kernel and a row kernel that re-read all B*A assignment words.  v5:
//
//   * LIST KERNEL (k7_loss_list_kernel): the assignment comes from k7_match_scatter as per-image lists
//     of the anchors a ground-truth box reached (a few per cent of A).  A thread per entry: positives go
//     to the image's queue (one returning atomic per warp, on the IMAGE's own counter - atomics on one
//     address serialise at ~7 ns each), ignored rows (losses.py:79-82) set a bit in the ignore bitmap.
//     Nothing walks the B*A anchors.  (k7_match's words instead: k7_loss_scan_kernel, one pass.)
//   * STREAM KERNEL (k7_loss_pipe_kernel).  One CTA per SM.  classifications is ONE flat array of
//     B*A*C floats cut into tiles of 16 x 512 floats; tiles ignore row AND image boundaries, so any
//     shape streams through 16-byte aligned bulk copies.  A producer warp drives a ring of
//     shared-memory stages with 1-D bulk async copies (TMA, SASS UBLKCP): tile + the ignore bits of
//     its rows in and - once the consumer warps have signalled the stage done - gradients out.
//     full[]/done[] mbarriers; no __syncthreads in the loop.
//   * FLAT PASS.  Consumer warp w owns floats [512w, 512w+512) of every tile: LDS.128 -> math ->
//     STS.128 on float4 number lane+32k (conflict free, balanced whatever C is).  EVERY element is
//     evaluated as "target 0" (negative / non-label class) with packed f32x2 arithmetic
//     (FFMA2/FMUL2, new on sm_100):
//         p = clamp(x); q = fl(1-p); pp = 1-q; nl = pp*S(pp) ~ -log(q)      (degree-5 minimax)
//         loss += p*p*nl                                  (x (1-alpha) once per image)
//         grad  = (p == x) ? (1-alpha)*scale * p * (2*nl + p/q) : 0          (MUFU.RCP)
//     The polynomial is valid for p <= 0.25; the pass tracks max(p), and in a slice that saw a larger
//     value the CONFIDENT elements alone are redone: their polynomial term leaves the sum, the exact
//     one (logf) enters, their gradient is overwritten.
//   * IGNORED ROWS never reach the arithmetic as such: before the flat pass the few elements of an
//     ignored row are overwritten IN THE TILE by a sentinel (0, or -1e30 for logits) whose clamped
//     probability differs from it - gradient 0 by torch.clamp's rule, loss term 1e-12.
//   * A slice that straddles an image boundary (B-1 of them per batch) or lies in a tile that
//     spans more than two images (tiny shapes) takes a generic per-element path.
//   * SERVE KERNEL (k7_loss_serve_kernel), after the stream: a thread per queued positive row:
//     smooth-L1 + d(reg), and the label element is corrected from "target 0" to "target 1" (loss
//     difference, gradient overwritten in dcls).
//   * Every contribution - the stream's per-(image, CTA) partials too - goes into two per-image
//     64-bit FIXED-POINT accumulators (2^-32 units, integer atomics): exact and independent of the
//     order in which blocks finish; the last block of the serve kernel turns them into the
//     per-image losses and the batch sums: bit-reproducible run to run, no reduction pass.
//   * The exchange of the batch sums between ranks rides on the serve kernel's last block (peer
//     memory over NVLink): fused (post + wait) or post-only with k7_peer_collect on another stream.
//
// Algorithmic bytes per image: forward A*(4C+16)+20M, forward+backward A*(8C+32)+20M.
#include <stdlib.h>
#include <string.h>

#include "k7_common.cuh"

#define K7_ROWS_THREADS 256
#define K7_FIX_ONE 4294967296.0                         // fixed-point unit of the per-image accumulators: 2^-32
#define K7_ROWS_PER 8                                      // anchors per thread and chunk
#define K7_ROWS_CHUNK (K7_ROWS_THREADS * K7_ROWS_PER)      // anchors per block and chunk

struct K7LossDev {
    float alpha, one_m_alpha;
    float beta, quad, shift;
    float div[4];
    float mean_div;
    int use_div;
    int yx_targets;
    int dreg_prezeroed;
    int from_logits;
    int B, A, C, M;
};

// where the row kernel takes the assignment from
struct K7AssignSrc {
    const int32_t* words;               // k7_match's packed state / row words, or
    const unsigned long long* best;     // k7_match_scatter's maxima (generation | IoU bits | ~row)
    unsigned int gen;
    float neg_thr, pos_thr;
    int32_t* words_out;                 // optional: the words, written on the way
    int consume;                        // best entries of this generation are zeroed once read (graph replay)
};

// The batch sums of the OTHER ranks, exchanged by the last block of the row kernel over peer memory
// (NVLink P2P stores into every rank's mailbox) instead of a separate all-reduce launch.
struct K7PeerDev {
    int world, rank;
    unsigned int seq;                  // call counter (>= 1), the same on every rank
    unsigned int* seq_counter;         // device word holding the number of calls so far, or NULL (then `seq` is used)
    int post_only;                     // post and return: the collection is k7_peer_collect's
    unsigned long long timeout_ns;
    float* mailbox[K7_MAX_PEERS];      // mailbox[p]: rank p's 2 * world slots of 8 floats {v0..v3, seq, -, -, -}
};

// per-image gradient scales, from the match stage's num_pos:
//   x: 1 / (max(num_pos,1) * mean_divisor)      (losses.py:118,168)
//   y: 1 / (4 * num_pos * mean_divisor) or 0    (losses.py:161,168)
__device__ __forceinline__ float2 k7_loss_scales(const int n, const float mean_div) {
    float2 s;
    s.x = 1.0f / (fmaxf((float)n, 1.0f) * mean_div);
    s.y = n > 0 ? 1.0f / (4.0f * (float)n * mean_div) : 0.0f;
    return s;
}

// ------------------------------------------------------------------ element math
__device__ __forceinline__ float k7_rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// sigmoid of two logits for the streaming pass: e = 2^(-z*log2 e) (MUFU.EX2), r = 1/(1+e) (MUFU.RCP);
// relative error ~3e-7, far inside the 1e-5 contract of the losses (the eval side, where the
// survivor set must be bit-exact, uses the precise form).
__device__ __forceinline__ float k7_ex2_approx(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float2 k7_sigmoid_pair(const float2 z) {
    const float2 t = __fmul2_rn(z, make_float2(-1.4426950408889634f, -1.4426950408889634f));
    float2 e;
    e.x = k7_ex2_approx(t.x);
    e.y = k7_ex2_approx(t.y);
    const float2 d = __fadd2_rn(e, make_float2(1.0f, 1.0f));
    float2 r;
    r.x = k7_rcp_approx(d.x);
    r.y = k7_rcp_approx(d.y);
    return r;
}
// The same with everything the gradient needs from ONE exponential and ONE reciprocal per element:
// ez = e^z (the exponent is capped so that it stays finite), r = 1/(1+ez) = 1 - sigma,
// sigma = ez*r, sigma' = sigma*r, and p/(1-p) = ez.
__device__ __forceinline__ float2 k7_sigmoid_pair_grad(const float2 z, float2& ez, float2& dsig) {
    float2 t = __fmul2_rn(z, make_float2(1.4426950408889634f, 1.4426950408889634f));
    t.x = fminf(t.x, 126.0f);
    t.y = fminf(t.y, 126.0f);
    ez.x = k7_ex2_approx(t.x);
    ez.y = k7_ex2_approx(t.y);
    const float2 d = __fadd2_rn(ez, make_float2(1.0f, 1.0f));
    float2 r;
    r.x = k7_rcp_approx(d.x);
    r.y = k7_rcp_approx(d.y);
    const float2 sg = __fmul2_rn(ez, r);
    dsig = __fmul2_rn(sg, r);
    return sg;
}

// Two "target 0" elements.  Returns their gradients (GRAD), accumulates p^2 * -log(1-p) into acc
// and max(p) into pmax.  LOGITS: the inputs are pre-sigmoid logits (the head's sigmoid is fused
// here) and the gradient is d/dz.
template <bool GRAD, bool LOGITS>
__device__ __forceinline__ float2 k7_neg_pair(const float2 xin, const float2 sco, float2& acc, float& pmax) {
    float2 dsig = make_float2(1.0f, 1.0f), ez = make_float2(0.0f, 0.0f);
    const float2 x = LOGITS ? (GRAD ? k7_sigmoid_pair_grad(xin, ez, dsig) : k7_sigmoid_pair(xin)) : xin;
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
    const float2 t = __fmul2_rn(p, p);
    acc = __ffma2_rn(t, nl, acc);
    float2 g = make_float2(0.0f, 0.0f);
    if (GRAD) {
        float2 u;                                                            // p / (1-p)
        if (LOGITS) {
            u = ez;   // exact identity for an unclamped p; a clamped p gets a zero gradient below anyway
        } else {
            float2 r;
            r.x = k7_rcp_approx(q.x);
            r.y = k7_rcp_approx(q.y);
            u = __fmul2_rn(p, r);
        }
        const float2 v = __ffma2_rn(make_float2(2.0f, 2.0f), nl, u);         // 2*nl + p/(1-p)
        float2 w = __fmul2_rn(p, sco);                                       // (1-alpha) * scale * p
        if (LOGITS) w = __fmul2_rn(w, dsig);                                 // chain rule through the sigmoid
        g = __fmul2_rn(w, v);
        // torch.clamp passes the gradient on the closed interval  <=>  clamp(x) == x
        g.x = (p.x == x.x) ? g.x : 0.0f;
        g.y = (p.y == x.y) ? g.y : 0.0f;
    }
    return g;
}

// One element as the flat pass sees it: the clamped probability p (for logits: through the same approximate sigmoid)
// and the polynomial loss term p^2 * (pp * S(pp)).  Used to take a CONFIDENT element (p > K7_POLY_MAX, outside the
// polynomial's range) out of the flat pass's sum again before it is added exactly.
template <bool GRAD, bool LOGITS>
__device__ __forceinline__ float k7_flat_term(const float xin, float& p_out) {
    float x = xin;
    if (LOGITS) {
        if (GRAD) {
            const float t = fminf(xin * 1.4426950408889634f, 126.0f);
            const float ez = k7_ex2_approx(t);
            x = ez * k7_rcp_approx(ez + 1.0f);
        } else {
            const float e = k7_ex2_approx(xin * -1.4426950408889634f);
            x = k7_rcp_approx(e + 1.0f);
        }
    }
    const float p = fminf(fmaxf(x, K7_P_LO), K7_P_HI);
    p_out = p;
    const float q = 1.0f - p, pp = 1.0f - q;
    return (p * p) * k7_neg_log1m_poly(pp);
}

template <bool GRAD, bool LOGITS>
__device__ __forceinline__ void k7_flat4(float4* __restrict__ p4, const float2 sco, float2& acc, float& pmax) {
    const float4 x = *p4;
    const float2 ga = k7_neg_pair<GRAD, LOGITS>(make_float2(x.x, x.y), sco, acc, pmax);
    const float2 gb = k7_neg_pair<GRAD, LOGITS>(make_float2(x.z, x.w), sco, acc, pmax);
    if (GRAD) *p4 = make_float4(ga.x, ga.y, gb.x, gb.y);
}

// The flat pass over the n4 float4s of a slice.
template <int TILE_F, bool GRAD, bool LOGITS>
__device__ __forceinline__ void k7_flat_pass(float* __restrict__ tile, const int n4, const int lane, const float sco1,
                                             float2& acc, float& pmax) {
    float4* t4 = reinterpret_cast<float4*>(tile);
    const float2 sco = make_float2(sco1, sco1);
    if (n4 == TILE_F / 4) {
#pragma unroll
        for (int k = 0; k < TILE_F / 128; ++k) k7_flat4<GRAD, LOGITS>(t4 + lane + 32 * k, sco, acc, pmax);
    } else {
        for (int f = lane; f < n4; f += 32) k7_flat4<GRAD, LOGITS>(t4 + f, sco, acc, pmax);
    }
}

// Exact scalar evaluation of one "target 0" element (any p): loss without the (1-alpha) factor,
// gradient without (1-alpha)*scale.  Used by the redo path and by the label patch.
// from_logits: x is a logit; the gradient is then d/dz.
__device__ __forceinline__ float k7_neg_exact(const float xin, float* grad, const bool from_logits = false) {
    const float x = from_logits ? k7_sigmoid_precise(xin) : xin;
    const float p = k7_clamp_p(x);
    const float q = 1.0f - p;
    const float nl = k7_neg_log1m(q);
    if (grad) {
        float g = (p == x) ? p * fmaf(2.0f, nl, p * k7_rcp_approx(q)) : 0.0f;
        if (from_logits) g *= x * (1.0f - x);
        *grad = g;
    }
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

// ------------------------------------------------------------------ the stream kernel
// One CTA per SM.  Warp CW is the PRODUCER: one elected lane cuts the flat batch into tiles of
// CW*SL floats, issues the TMA loads into a ring of NS stages, and, once the CW consumer warps have
// signalled a stage done, the TMA store of the gradients.  Warps 0..CW-1 are CONSUMERS: warp w owns
// floats [SL*w, SL*w+SL) of every tile (SL = 512 or 1024).  full[s] / done[s] are mbarriers; there
// is no __syncthreads after the prologue.
#define K7_PIPE_MAX_CW 16

// One contiguous stream of a launch: the [B,A,C] tensor, or one pyramid level of the per-level layout.
struct K7PipeSeg {
    const float* cls;
    float* dcls;
    const uint32_t* bits;    // one bit per ROW of the stream, set = ignored (losses.py:79-82); image b owns words [b*W, (b+1)*W)
    unsigned long long N;    // B * AC elements
    unsigned long long N4;   // N & ~3: the part moved by 16-byte bulk copies
    unsigned long long AC;   // elements per image
    uint32_t A;              // rows per image (a row = C elements; per-level streams run with C == 1: a row per element)
    uint32_t W;              // bitmap words per image, K7_IGNORE_WORDS(A)
    uint32_t tile0;          // first tile of the stream in the launch's tile order
    uint32_t pad;
};
struct K7PipeSegs {
    int n, pad;
    K7PipeSeg s[K7_MAX_LEVELS];
};

#define K7_PIPE_SCO_MAX 4096   // images whose gradient scale fits the shared-memory table
struct K7PipeGeom {
    uint32_t tile_f;         // floats per tile = cw * slice
    uint32_t slice;          // floats per consumer warp and tile: 512 or 1024
    uint32_t num_tiles;      // sum over the streams of ceil(N4 / tile_f)
    K7FastDiv div_c;         // / C for tile-local indices
    uint32_t cta_stride, cta_off;   // (unused since the stream folds into the per-image accumulators)
    uint32_t side_words;     // words reserved per ignore-bit part (two parts per stage: image b0 and image b0 + 1)
    uint32_t stage_stride;
    uint32_t off_desc, off_bars, off_sco;   // off_sco: float[B] gradient scale per image (0: not tabulated)
    int stages, cw, grid;
    size_t smem_bytes;
};

struct K7TileDesc {   // written by the producer next to each stage, 48 bytes
    uint32_t n, b0, split, split2;     // floats in the tile | image of float 0 | floats that belong to image b0 | ... up to the end of image b0+1
    uint32_t off_lo, off_hi, seg, pad; // flat index of float 0 in its stream | the stream
    uint32_t rem, lead, kA, a0;        // position of float 0 in its row | bit of tile row 0 in part 0 | tile row where image b0+1 starts | row (in image b0) of tile row 0
};

template <bool GRAD, int SL, bool LOGITS>
__global__ void __launch_bounds__((K7_PIPE_MAX_CW + 1) * 32, 1)
k7_loss_pipe_kernel(const __grid_constant__ K7PipeSegs SG, const int32_t* __restrict__ num_pos,
                    const unsigned char* __restrict__ img_valid, const __grid_constant__ K7LossDev P,
                    const __grid_constant__ K7PipeGeom G, float* __restrict__ partials, long long* __restrict__ img_acc) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int NS = G.stages, CW = G.cw;
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
    k7_pdl_trigger();   // programmatic dependent launch (no-ops in a plain launch): the serve kernel's CTAs may take
    k7_pdl_wait();      // their places; the list kernel's outputs (num_pos, img_valid, the bitmap) are complete
    // The gradient scale of every image, (1-alpha) / (max(num_pos,1) * mean_divisor) - 0 for an image
    // without ground truth, which the reference skips (losses.py:52-60) -, tabulated once: a warp that
    // enters a new image must not wait for global memory (all 16 warps of a CTA change image at about
    // the same tile: the whole CTA would stall, 64 times per kernel).
    const float* s_sco = (GRAD && G.off_sco) ? reinterpret_cast<const float*>(smem + G.off_sco) : nullptr;
    if (GRAD && G.off_sco) {
        float* t = reinterpret_cast<float*>(smem + G.off_sco);
        for (int b = threadIdx.x; b < P.B; b += blockDim.x)
            t[b] = img_valid[b] ? P.one_m_alpha * k7_loss_scales(__ldg(num_pos + b), P.mean_div).x : 0.0f;
    }
    __syncthreads();
    if (blockIdx.x >= G.num_tiles && blockIdx.x != 0) return;
    const uint32_t ntl = blockIdx.x < G.num_tiles ? (G.num_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;   // tiles of this CTA

    if (warp == CW) {
        // ================================================================ producer
        if (lane != 0) return;
        // The descriptor of the NEXT tile is worked out (64-bit divisions) while the producer would
        // otherwise wait for a stage: a tile lasts ~0.8 us, the arithmetic must stay off the path
        // between "stage free" and "copy issued".
        K7TileDesc nd;
        const float* n_src = nullptr;
        const uint32_t *n_bits0 = nullptr, *n_bits1 = nullptr;
        uint32_t n_words0 = 0u, n_words1 = 0u;
        auto prepare = [&](const uint32_t j) {
            const uint32_t g = blockIdx.x + j * gridDim.x;
            int sg = 0;
            for (int i = 1; i < SG.n; ++i)
                if (g >= SG.s[i].tile0) sg = i;
            const K7PipeSeg& Q = SG.s[sg];
            const unsigned long long off = (unsigned long long)(g - Q.tile0) * G.tile_f;
            const unsigned long long left = Q.N4 - off;
            K7TileDesc d;
            d.n = left < G.tile_f ? (uint32_t)left : G.tile_f;
            d.b0 = (uint32_t)(off / Q.AC);
            const unsigned long long in_img = (unsigned long long)(d.b0 + 1u) * Q.AC - off;
            d.split = in_img < d.n ? (uint32_t)in_img : d.n;
            const unsigned long long in_img2 = in_img + Q.AC;
            d.split2 = in_img2 < d.n ? (uint32_t)in_img2 : d.n;
            d.off_lo = (uint32_t)off;
            d.off_hi = (uint32_t)(off >> 32);
            d.seg = (uint32_t)sg;
            d.pad = 0u;
            // the ignore bits of the rows touching the tile: part 0 = rows of image b0, part 1 = rows of image b0 + 1
            const uint32_t A = Q.A;
            const unsigned long long r0 = off / (uint32_t)P.C;   // first row (batch-global) touching the tile
            d.rem = (uint32_t)(off - r0 * (uint32_t)P.C);
            d.a0 = (uint32_t)(r0 - (unsigned long long)d.b0 * A);
            const uint32_t nrows = k7_div(d.rem + d.n - 1u, G.div_c) + 1u;
            d.kA = A - d.a0;
            const uint32_t w0 = (d.a0 >> 7) << 2;   // 16-byte aligned first word of part 0
            d.lead = d.a0 - (w0 << 5);
            const uint32_t rows0 = nrows < d.kA ? nrows : d.kA;
            n_words0 = ((d.lead + rows0 + 127u) >> 7) << 2;
            n_words1 = 0u;
            if (nrows > d.kA && d.b0 + 1u < (uint32_t)P.B) {
                const uint32_t rows1 = min(nrows - d.kA, A);
                n_words1 = ((rows1 + 127u) >> 7) << 2;
            }
            nd = d;
            n_src = Q.cls + off;
            n_bits0 = Q.bits + (size_t)d.b0 * Q.W + w0;
            n_bits1 = Q.bits + (size_t)(d.b0 + 1u) * Q.W;
        };
        auto issue = [&](const uint32_t j) {   // the prepared tile is tile j
            const int s = (int)(j % (uint32_t)NS);
            unsigned char* stage = smem + (size_t)s * G.stage_stride;
            desc[s] = nd;
            uint32_t* side = reinterpret_cast<uint32_t*>(stage + (size_t)G.tile_f * 4);
            k7_mbar_expect_tx(full + s, nd.n * 4u + (n_words0 + n_words1) * 4u);   // release: orders the descriptor
            k7_bulk_g2s(stage, n_src, nd.n * 4u, full + s);
            k7_bulk_g2s(side, n_bits0, n_words0 * 4u, full + s);
            if (n_words1) k7_bulk_g2s(side + G.side_words, n_bits1, n_words1 * 4u, full + s);
        };
        // with gradients one stage is always draining its store
        const uint32_t ahead = (uint32_t)(GRAD ? NS - 1 : NS);
        uint32_t jl = 0;
        if (ntl > 0) prepare(0);
        while (jl < ntl && jl < ahead) {
            issue(jl++);
            if (jl < ntl) prepare(jl);
        }
        int s = 0;
        uint32_t phase = 0;
        for (uint32_t j = 0; j < ntl; ++j) {
            k7_mbar_wait(done + s, phase);
            if (GRAD) {
                const K7TileDesc d = desc[s];
                const unsigned long long off = ((unsigned long long)d.off_hi << 32) | d.off_lo;
                k7_bulk_s2g(SG.s[d.seg].dcls + off, smem + (size_t)s * G.stage_stride, d.n * 4u);
                k7_bulk_commit();
                if (jl < ntl) {
                    k7_bulk_wait_read<1>();   // the store of tile j-1 has left its stage
                    issue(jl++);
                    if (jl < ntl) prepare(jl);
                }
            } else if (jl < ntl) {
                issue(jl++);
                if (jl < ntl) prepare(jl);
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
    float* my_partials = partials + (size_t)blockIdx.x * CW + warp;      // + b * (gridDim.x * CW)
    const size_t pstride = (size_t)gridDim.x * CW;
    for (int b = lane; b < P.B; b += 32) my_partials[(size_t)b * pstride] = 0.0f;
    __syncwarp();

    const float oma = P.one_m_alpha;
    float2 acc = make_float2(0.0f, 0.0f);   // sum p^2 * -log(1-p) of the current image (this lane), polynomial path
    float cslow = 0.0f;                     // the same from the exact paths
    int cur_b = -1;
    float sco = 0.0f;
    auto flush = [&]() {
        float c = oma * ((acc.x + acc.y) + cslow);
        c = k7_warp_sum(c);
        if (lane == 0) atomicAdd(my_partials + (size_t)cur_b * pstride, c);   // RED: no round trip; one adder per slot, in program order
        acc = make_float2(0.0f, 0.0f);
        cslow = 0.0f;
    };
    // Within a stream a warp meets the images in increasing order; a later stream (the next pyramid
    // level) starts over, so a slot accumulates: always by lane 0 of this warp, in program order.
    auto enter = [&](const int b) {
        if (b == cur_b) return;
        if (cur_b >= 0) flush();
        cur_b = b;
        if (GRAD)
            sco = s_sco ? s_sco[b] : (img_valid[b] ? oma * k7_loss_scales(__ldg(num_pos + b), P.mean_div).x : 0.0f);
    };
    auto exact = [&](const float x, const bool ign, float* gout) {
        float gr;
        const float l = k7_neg_exact(x, &gr, LOGITS);
        if (!ign) cslow += l;
        if (GRAD) *gout = ign ? 0.0f : sco * gr;
    };
    // An element of an IGNORED row (losses.py:79-82) is overwritten in the shared-memory tile, before the flat
    // pass, by a value whose clamped probability differs from it: clamp(0) = 1e-4 != 0 gives gradient 0
    // (torch.clamp's rule) and a loss term p^2 * -log(1-p) = 1e-12 - below half an ulp of any sum it is added
    // to, < 1e-9 relative in total.  The flat pass itself knows nothing about rows: no mask, no second variant.
    const float SENT = LOGITS ? -1.0e30f : 0.0f;
    const int C = P.C;
    const int i_lo = warp * SL;

    const uint32_t full_u32 = k7_smem_u32(full), done_u32 = k7_smem_u32(done);
    int s = -1;
    uint32_t phase = 1u;
    for (uint32_t j = 0; j < ntl; ++j) {
        if (++s == NS) s = 0;
        if (s == 0) phase ^= 1u;
        unsigned char* stage = smem + (size_t)s * G.stage_stride;
        k7_mbar_wait_u32(full_u32 + 8u * s, phase);
        const uint4 d0 = *reinterpret_cast<const uint4*>(desc + s);   // n, b0, split, split2
        const int n = (int)d0.x;
        if (i_lo < n) {
            const int i_hi = min(n, i_lo + SL);
            const int len = i_hi - i_lo, n4 = len >> 2;
            float* tile = reinterpret_cast<float*>(stage);
            float* slice = tile + i_lo;
            int img = -1;   // which image does the slice lie in?
            if ((uint32_t)i_hi <= d0.z) img = (int)d0.y;
            else if ((uint32_t)i_lo >= d0.z && (uint32_t)i_hi <= d0.w) img = (int)d0.y + 1;
            const uint4 d2 = *(reinterpret_cast<const uint4*>(desc + s) + 2);   // rem, lead, kA, a0
            const int rem = (int)d2.x;
            int bitbase = 0;   // bit position of tile row 0 in the part of the slice's image
            if (img == (int)d0.y) bitbase = (int)d2.y;
            else if (img >= 0) bitbase = (int)(G.side_words * 32u) - (int)d2.z;   // part 1 starts at tile row kA
            if (img >= 0) {
                enter(img);
                const uint32_t* side = reinterpret_cast<const uint32_t*>(stage + (size_t)G.tile_f * 4);
                const int k_lo = (int)k7_div((uint32_t)(rem + i_lo), G.div_c);
                const int k_hi = (int)k7_div((uint32_t)(rem + i_hi - 1), G.div_c);
                // ---- ignored rows of the slice: their elements become SENT
                bool masked = false;
                if (C == 1) {
                    // a row is one element: 32 consecutive bits of the bitmap per 32 elements
                    for (int k0 = k_lo + 32 * lane; k0 <= k_hi; k0 += 1024) {
                        const int p = bitbase + k0;
                        uint32_t w = __funnelshift_r(side[p >> 5], side[(p >> 5) + 1], (uint32_t)p & 31u);
                        const int nv = k_hi - k0 + 1;
                        if (nv < 32) w &= (1u << nv) - 1u;
                        if (w) masked = true;
                        while (w) {
                            const int j = __ffs(w) - 1;
                            w &= w - 1u;
                            slice[k0 - k_lo + j] = SENT;
                        }
                    }
                    masked = __any_sync(0xffffffffu, masked);
                } else {
                    for (int k0 = k_lo; k0 <= k_hi; k0 += 32) {   // warp-uniform window reads
                        const int p = bitbase + k0;
                        uint32_t win = __funnelshift_r(side[p >> 5], side[(p >> 5) + 1], (uint32_t)p & 31u);
                        const int nv = k_hi - k0 + 1;
                        if (nv < 32) win &= (1u << nv) - 1u;
                        if (win == 0u) continue;
                        masked = true;
                        while (win) {
                            const int kk = k0 + __ffs(win) - 1;
                            win &= win - 1u;
                            const int e_lo = kk * C - rem - i_lo;
                            for (int e = max(e_lo, 0) + lane; e < min(e_lo + C, len); e += 32) slice[e] = SENT;
                        }
                    }
                }
                if (masked) __syncwarp();
                float pmax = 0.0f;
                k7_flat_pass<SL, GRAD, LOGITS>(slice, n4, lane, sco, acc, pmax);
                if (__any_sync(0xffffffffu, pmax > K7_POLY_MAX)) {
                    // Confident predictions (p > 0.25: outside the polynomial's range) in this slice.  Only THOSE elements
                    // are redone: their polynomial term leaves the sum again and the exact one (logf) enters, their
                    // gradient is overwritten.  With gradients the tile already holds gradients, so the inputs come from
                    // global memory again (L2: the tile was loaded a moment ago).  (Redoing the whole slice exactly, as
                    // before, cost 7 x the flat pass per slice: at C3 - one slice in five - it was half of the kernel.)
                    __syncwarp();
                    const uint4 d1 = *(reinterpret_cast<const uint4*>(desc + s) + 1);   // off_lo, off_hi, seg
                    const float* src = SG.s[d1.z].cls + ((((unsigned long long)d1.y) << 32) | d1.x) + i_lo;
                    float fix = 0.0f;
                    for (int i0 = lane; i0 < len; i0 += 128) {
                        float xv[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int i = i0 + 32 * u;
                            xv[u] = i < len ? (GRAD ? __ldg(src + i) : slice[i]) : 0.0f;   // (0: never confident)
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int i = i0 + 32 * u;
                            float pf;
                            const float term = k7_flat_term<GRAD, LOGITS>(xv[u], pf);
                            if (!(pf > K7_POLY_MAX) || i >= len) continue;
                            if (GRAD && masked) {   // an ignored row: the tile holds its zero gradient, its term was ~0
                                const int p = bitbase + (int)k7_div((uint32_t)(rem + i_lo + i), G.div_c);
                                if ((side[p >> 5] >> (p & 31)) & 1u) continue;
                            }
                            fix -= term;
                            exact(xv[u], false, slice + i);
                        }
                    }
                    acc.x += fix;
                }
            } else {
                // ---- generic path: the slice straddles an image boundary (or the tile spans > 2 images)
                const uint4 d1 = *(reinterpret_cast<const uint4*>(desc + s) + 1);
                const unsigned long long off = (((unsigned long long)d1.y) << 32) | d1.x;
                const K7PipeSeg& Q = SG.s[d1.z];
                const uint32_t A = Q.A;
                const unsigned long long r0 = (unsigned long long)d0.y * A + d2.w;
                uint32_t i = (uint32_t)i_lo;
                while (i < (uint32_t)i_hi) {   // warp-uniform: one iteration per image segment
                    const uint32_t b = (uint32_t)((off + i) / Q.AC);
                    const unsigned long long img_end = (unsigned long long)(b + 1u) * Q.AC - off;
                    const uint32_t seg_end = img_end < (unsigned long long)i_hi ? (uint32_t)img_end : (uint32_t)i_hi;
                    enter((int)b);
                    const unsigned long long abase = (unsigned long long)b * A;
                    for (uint32_t t = i + lane; t < seg_end; t += 32) {
                        const uint32_t k = k7_div((uint32_t)rem + t, G.div_c);
                        const uint32_t a = (uint32_t)(r0 + k - abase);
                        const bool ign = (__ldg(Q.bits + (size_t)b * Q.W + (a >> 5)) >> (a & 31u)) & 1u;
                        exact(tile[t], ign, tile + t);
                    }
                    i = seg_end;
                }
            }
        }
        if (GRAD) k7_fence_proxy_async();
        __syncwarp();
        if (lane == 0) k7_mbar_arrive_u32(done_u32 + 8u * s);
    }

    // the last N % 4 elements cannot be bulk-copied: CTA 0, warp 0 serves them from global memory
    if (blockIdx.x == 0 && warp == 0) {
        for (int sg = 0; sg < SG.n; ++sg) {
            const K7PipeSeg& Q = SG.s[sg];
            for (unsigned long long e = Q.N4; e < Q.N; ++e) {
                const uint32_t b = (uint32_t)(e / Q.AC);
                enter((int)b);
                if (lane == 0) {
                    const unsigned long long r = e / (uint32_t)C;
                    const uint32_t a = (uint32_t)(r - (unsigned long long)b * Q.A);
                    const bool ign = (__ldg(Q.bits + (size_t)b * Q.W + (a >> 5)) >> (a & 31u)) & 1u;
                    float g = 0.0f;
                    exact(__ldg(Q.cls + e), ign, &g);
                    if (GRAD) Q.dcls[e] = g;
                }
            }
        }
    }
    if (cur_b >= 0) flush();

    // ---- fold the CTA's (image, warp) slots into the image's fixed-point accumulator (2^-32 units, integer
    // atomics: exact, whatever the order in which CTAs finish - nobody has to sum 148 partials per image later)
    asm volatile("bar.sync 1, %0;" ::"r"(CW * 32) : "memory");
    // a lane per image (the CW slots of an image are contiguous): one trip for up to CW * 32 images
    for (int b = warp * 32 + lane; b < P.B; b += CW * 32) {
        const float* sl = partials + (size_t)b * pstride + (size_t)blockIdx.x * CW;
        float v = 0.0f;
        for (int w = 0; w < CW; ++w) v += __ldcg(sl + w);   // fixed order
        if (v != 0.0f)
            atomicAdd(reinterpret_cast<unsigned long long*>(img_acc + 2 * b), (unsigned long long)__double2ll_rn((double)v * K7_FIX_ONE));
    }
}

// ------------------------------------------------------------------ head outputs per level
// levels[l] = [B, K*C, S_l] (NCHW, S_l = fh*fw cells, channel = in-cell anchor k * C + class c): the
// element of anchor a = offset[l] + s*K + k and class c sits at ((b*K*C + k*C + c) * S_l + s).
// Every element is a "target 0" element whatever the layout, so the stream kernel runs over the
// levels as ONE launch with a table of streams; the row kernel addresses the label element and
// the elements of an ignored row through the level table below.
struct K7LevelTable {
    const float* cls[K7_MAX_LEVELS];
    float* dcls[K7_MAX_LEVELS];
    int S[K7_MAX_LEVELS];
};

// Seen as a flat stream every element of a per-level tensor is its own "row": the stream kernel runs on a
// level in its C == 1 mode with ONE IGNORE BIT PER ELEMENT, laid out like the tensor: bit (b, ch, s) = the
// anchor's bit.  k7_bits_expand_kernel writes those words (1/32 of the tensor's bytes) from the
// anchor-ordered bitmap.
struct K7ExpandTable {
    uint32_t* lbits[K7_MAX_LEVELS];
    int level_off[K7_MAX_LEVELS], S[K7_MAX_LEVELS], Wl[K7_MAX_LEVELS];
    int blk_off[K7_MAX_LEVELS + 1];   // first block of each level; [K7_MAX_LEVELS] = grid.x
};

// All levels in one launch, grid (blocks, B).  One thread per 32-cell word of an (image, in-cell
// anchor) plane; the word is the same for all C classes of that anchor.  S % 32 == 0: whole words,
// written C times.  Otherwise (small levels) one thread per output word gathers its 32 elements
// individually.
__global__ void __launch_bounds__(256)
k7_bits_expand_kernel(const uint32_t* __restrict__ abits, int W, const __grid_constant__ K7ExpandTable T, int K, int C) {
    const int b = blockIdx.y;
    int l = 0;
#pragma unroll
    for (int i = 1; i < K7_MAX_LEVELS; ++i)
        if ((int)blockIdx.x >= T.blk_off[i]) l = i;
    const int S = T.S[l], level_off = T.level_off[l];
    const int blk = (int)blockIdx.x - T.blk_off[l];
    const uint32_t* ab = abits + (size_t)b * W;
    uint32_t* lb = T.lbits[l] + (size_t)b * T.Wl[l];
    if ((S & 31) == 0) {
        const int spw = S >> 5;                       // words per plane
        const int t = blk * 256 + (int)threadIdx.x;
        if (t >= K * spw) return;
        const int k = t / spw, sw = t - k * spw;
        uint32_t word = 0u;
        int a = level_off + (sw << 5) * K + k;
#pragma unroll 8
        for (int j = 0; j < 32; ++j, a += K) word |= ((__ldg(ab + (a >> 5)) >> (a & 31)) & 1u) << j;
        for (int c = 0; c < C; ++c) lb[(size_t)(k * C + c) * spw + sw] = word;
    } else {
        const int nel = K * C * S;                    // < 2^31 (checked by the plan)
        const int w = blk * 256 + (int)threadIdx.x;
        if (w >= (nel + 31) >> 5) return;
        const int e0 = w << 5;
        int ch = e0 / S, s = e0 - ch * S;             // plane (k*C + c) and cell of the word's first element
        int k = ch / C, c = ch - k * C;
        const int nb = min(32, nel - e0);
        uint32_t word = 0u;
        for (int j = 0; j < nb; ++j) {
            const int a = level_off + s * K + k;
            word |= ((__ldg(ab + (a >> 5)) >> (a & 31)) & 1u) << j;
            if (++s == S) {                           // next plane: same anchor plane until the classes run out
                s = 0;
                if (++c == C) {
                    c = 0;
                    ++k;
                }
            }
        }
        lb[w] = word;
    }
}

// ------------------------------------------------------------------ the row kernels
__device__ __forceinline__ double k7_warp_sum_f64(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ long long k7_warp_sum_i64(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Sum of (c, r) over the block, result valid in thread 0.  Fixed order.  s_red: 2 * warps doubles.
__device__ __forceinline__ void k7_block_sum2_f64(double& c, double& r, double* s_red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int NW = K7_ROWS_THREADS / 32;
    c = k7_warp_sum_f64(c);
    r = k7_warp_sum_f64(r);
    __syncthreads();   // s_red may still be read from a previous use
    if (lane == 0) {
        s_red[warp] = c;
        s_red[NW + warp] = r;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        c = 0.0;
        r = 0.0;
#pragma unroll
        for (int w = 0; w < NW; ++w) {
            c += s_red[w];
            r += s_red[NW + w];
        }
    }
}

// SCAN KERNEL, before the stream.  Work item = a chunk of 2048 consecutive anchors of one image;
// every thread reads 8 assignment entries (coalesced: entry j of thread t is anchor a0 + 256 j + t),
// zeroes d(reg) of the rows that are not positive, counts the image's positives (num_pos: the
// stream needs it for the gradient scale) and appends the rows that are not plain negatives to a
// batch-wide queue in global memory - positive rows from the front, ignored rows (losses.py:79-82)
// from the back; one block-level prefix and ONE atomicAdd per block and queue end.  The order of
// the queue depends on the order in which blocks finish; everything computed from it is
// accumulated with integer atomics, so results do not.
//   queue entry: x = anchor (image-local), y = image << 16 | annotation row
struct K7RowQueues {
    uint2* q;                  // [B][A] entries (anchor, annotation row): the positive rows of image b are q[b*A .. b*A + count_b)
    unsigned int* count;       // count[b * K7_BEST_CNT_STRIDE]: positives of image b (one counter per image, apart: atomics
                               // on one address serialise)
    unsigned long long cap;    // B*A
    unsigned char* img_valid;  // [B] 1: the image has an annotation row (losses.py:49-60)
    uint32_t* bits;            // [B][W] one bit per anchor, set = ignored row (losses.py:79-82): what the stream reads
    int W;                     // K7_IGNORE_WORDS(A)
};

template <bool GRAD, bool PACKED>
__global__ void __launch_bounds__(K7_ROWS_THREADS)
k7_loss_scan_kernel(const float* __restrict__ annotations, const __grid_constant__ K7AssignSrc SRC,
                    const __grid_constant__ K7LossDev P, const __grid_constant__ K7RowQueues Q,
                    int32_t* __restrict__ num_pos, float4* __restrict__ dreg) {
    constexpr int NW = K7_ROWS_THREADS / 32;
    __shared__ int s_wp[NW];
    __shared__ unsigned int s_base[2];
    const int A = P.A;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cpi = (A + K7_ROWS_CHUNK - 1) / K7_ROWS_CHUNK;   // chunks per image
    const long long total = (long long)P.B * cpi;

    for (long long ch = blockIdx.x; ch < total; ch += gridDim.x) {   // block-uniform
        const int b = (int)(ch / cpi);
        const int a0 = (int)(ch - (long long)b * cpi) * K7_ROWS_CHUNK;
        // the assignment entries first: their latency overlaps the look at the annotations
        unsigned long long e[K7_ROWS_PER];
        int w[K7_ROWS_PER];
#pragma unroll
        for (int j = 0; j < K7_ROWS_PER; ++j) {
            const int a = a0 + j * K7_ROWS_THREADS + threadIdx.x;
            if (PACKED) e[j] = (a < A) ? __ldcg(SRC.best + (size_t)b * A + a) : 0ull;
            else w[j] = (a < A) ? __ldg(SRC.words + (size_t)b * A + a) : 0;
        }
        bool v = false;
        for (int m = threadIdx.x; m < P.M; m += K7_ROWS_THREADS) v |= __ldg(annotations + ((size_t)b * P.M + m) * 5 + 4) != -1.0f;
        const bool any_valid = __syncthreads_or(v ? 1 : 0) != 0;   // also: s_wp / s_wi / s_base of the previous chunk are consumed
        if (a0 == 0 && threadIdx.x == 0) Q.img_valid[b] = any_valid ? 1 : 0;

        unsigned int pm = 0u;   // bit j: entry j is a positive
        int rw[K7_ROWS_PER];
#pragma unroll
        for (int j = 0; j < K7_ROWS_PER; ++j) {
            const int a = a0 + j * K7_ROWS_THREADS + threadIdx.x;
            int st = K7_STATE_NEG;
            rw[j] = 0;
            if (PACKED) {
                if (K7_BEST_GEN(e[j]) == SRC.gen) {
                    if (SRC.consume) const_cast<unsigned long long*>(SRC.best)[(size_t)b * A + a] = 0ull;
                    const float iou = __uint_as_float(K7_BEST_IOU_BITS(e[j]));
                    st = iou >= SRC.pos_thr ? K7_STATE_POS : (iou < SRC.neg_thr ? K7_STATE_NEG : K7_STATE_IGNORE);
                    rw[j] = K7_BEST_ROW(e[j]);
                }
            } else {
                st = K7_ASSIGN_STATE(w[j]);
                rw[j] = K7_ASSIGN_ROW(w[j]);
            }
            if (a < A && !any_valid) {   // the whole image is skipped (losses.py:52-60): nothing to patch, no gradient
                st = K7_STATE_IGNORE;
                rw[j] = 0;
            }
            // the stream's bitmap: 32 consecutive anchors per warp and entry -> one word (a0 is a multiple of 2048)
            const unsigned int bal = __ballot_sync(0xffffffffu, a < A && st == K7_STATE_IGNORE && any_valid);
            const int aw = a0 + j * K7_ROWS_THREADS + warp * 32;
            if (lane == 0 && aw < A) Q.bits[(size_t)b * Q.W + (aw >> 5)] = bal;
            if (a >= A) continue;
            const size_t R = (size_t)b * A + a;
            if (SRC.words_out) SRC.words_out[R] = (rw[j] << 2) | st;
            if (GRAD && !P.dreg_prezeroed && st != K7_STATE_POS) dreg[R] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (any_valid) pm |= (st == K7_STATE_POS ? 1u : 0u) << j;
        }
        // block-level exclusive prefix of the positives per thread
        const int cp = __popc(pm);
        int ip = cp;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int tp = __shfl_up_sync(0xffffffffu, ip, o);
            if (lane >= o) ip += tp;
        }
        if (lane == 31) s_wp[warp] = ip;
        __syncthreads();
        int op = ip - cp, tp = 0;
#pragma unroll
        for (int k = 0; k < NW; ++k) {
            if (k < warp) op += s_wp[k];
            tp += s_wp[k];
        }
        if (threadIdx.x == 0) {
            s_base[0] = tp ? atomicAdd(Q.count + (size_t)b * K7_BEST_CNT_STRIDE, (unsigned int)tp) : 0u;
            if (tp) atomicAdd(num_pos + b, tp);   // integer: order independent, exact
        }
        __syncthreads();
        if (pm) {
            const size_t bp = (size_t)b * A + s_base[0] + (unsigned int)op;
            int kp = 0;
#pragma unroll
            for (int j = 0; j < K7_ROWS_PER; ++j) {
                const unsigned int a = (unsigned int)(a0 + j * K7_ROWS_THREADS + threadIdx.x);
                if ((pm >> j) & 1u) Q.q[bp + kp++] = make_uint2(a, (unsigned int)rw[j]);
            }
        }
    }
}

// LIST KERNEL - the scan's job when the assignment comes from k7_match_scatter: that kernel leaves, behind
// its array of maxima, the list of anchors it TOUCHED (IoU_max >= neg_thr: every row that is not a plain
// negative, a few per cent of A), so nothing has to walk the B*A entries.  A thread per list entry reads the
// final maximum; a positive is counted for its image and appended to the queue (one atomicAdd per warp), an
// ignored row sets its bit in the stream's bitmap (zeroed at the start of the call).  No block-wide step.
// Also: img_valid[b].
#define K7_LIST_PER 4   // entries per thread and trip: four independent load chains in flight
__global__ void __launch_bounds__(K7_ROWS_THREADS)
k7_loss_list_kernel(const float* __restrict__ annotations, const __grid_constant__ K7AssignSrc SRC,
                    const __grid_constant__ K7LossDev P, const __grid_constant__ K7RowQueues Q,
                    int32_t* __restrict__ num_pos, const unsigned int* __restrict__ tcount,
                    const unsigned int* __restrict__ tlist) {
    const int lane = threadIdx.x & 31;
    const unsigned int b = blockIdx.y;   // grid (blocks per image, B)
    const unsigned int gw = (blockIdx.x * K7_ROWS_THREADS + threadIdx.x) >> 5, nw = (gridDim.x * K7_ROWS_THREADS) >> 5;
    if (gw == 0) {
        bool v = false;
        for (int m = lane; m < P.M; m += 32) v |= __ldg(annotations + ((size_t)b * P.M + m) * 5 + 4) != -1.0f;
        v = __any_sync(0xffffffffu, v);
        if (lane == 0) Q.img_valid[b] = v ? 1 : 0;
    }
    const unsigned int n = __ldcg(tcount + (size_t)b * K7_BEST_CNT_STRIDE);
    const unsigned int* list = tlist + (size_t)b * P.A;
    const unsigned long long* best = SRC.best + (size_t)b * P.A;
    for (unsigned int i0 = gw * (32 * K7_LIST_PER); i0 < n; i0 += nw * (32 * K7_LIST_PER)) {   // warp-uniform
        unsigned int an[K7_LIST_PER];
        unsigned long long e[K7_LIST_PER];
#pragma unroll
        for (int u = 0; u < K7_LIST_PER; ++u) {
            const unsigned int i = i0 + 32 * u + lane;
            an[u] = i < n ? __ldcg(list + i) : 0u;
        }
#pragma unroll
        for (int u = 0; u < K7_LIST_PER; ++u) {
            const unsigned int i = i0 + 32 * u + lane;
            e[u] = i < n ? __ldcg(best + an[u]) : 0ull;
        }
        bool pos[K7_LIST_PER];
        int npos = 0;
#pragma unroll
        for (int u = 0; u < K7_LIST_PER; ++u) {
            const unsigned int a = an[u];
            bool ign = false;
            pos[u] = false;
            if (K7_BEST_GEN(e[u]) == SRC.gen) {   // (an entry beyond n reads 0: generation 0 is never used)
                if (SRC.consume) const_cast<unsigned long long*>(best)[a] = 0ull;
                const float iou = __uint_as_float(K7_BEST_IOU_BITS(e[u]));
                pos[u] = iou >= SRC.pos_thr;
                ign = !pos[u] && !(iou < SRC.neg_thr);
            }
            if (ign) atomicOr(Q.bits + (size_t)b * Q.W + (a >> 5), 1u << (a & 31u));
            npos += __popc(__ballot_sync(0xffffffffu, pos[u]));
        }
        if (npos == 0) continue;   // warp-uniform
        unsigned int base = 0u;    // ONE atomic per warp and trip, on the image's own counter
        if (lane == 0) {
            base = atomicAdd(Q.count + (size_t)b * K7_BEST_CNT_STRIDE, (unsigned int)npos);
            atomicAdd(num_pos + b, npos);
        }
        base = __shfl_sync(0xffffffffu, base, 0);
#pragma unroll
        for (int u = 0; u < K7_LIST_PER; ++u) {
            const unsigned int pm = __ballot_sync(0xffffffffu, pos[u]);
            if (pos[u]) Q.q[(size_t)b * P.A + base + (unsigned int)__popc(pm & ((1u << lane) - 1u))] = make_uint2(an[u], (unsigned int)K7_BEST_ROW(e[u]));
            base += (unsigned int)__popc(pm);
        }
    }
}

// SERVE KERNEL, after the stream: the queue of positive rows, densely, a thread each: smooth-L1 + d(reg)
// (losses.py:122-166) and the correction of the label element from "target 0" (what the stream computed)
// to "target 1" (losses.py:94-112).  (Ignored rows never reach it: the stream skips them by the bitmap.)
// Contributions are added to the image's two accumulators as 2^-32 fixed-point integers (integer
// atomics: exact, independent of the order) - the stream's CTAs added theirs the same way.  The last
// block to finish turns the accumulators into (cls_loss_i, reg_loss_i), then the batch sums in image
// order.  No float atomics anywhere: bit-reproducible run to run.
__device__ __forceinline__ void k7_acc_add(long long* __restrict__ acc, int b, int which, double v, int lane) {
    // one atomic per image present in the warp (neighbours in the queue mostly share their image; b < 0: no row)
    const long long iv = b >= 0 ? __double2ll_rn(v * K7_FIX_ONE) : 0ll;
    unsigned int left = __ballot_sync(0xffffffffu, b >= 0);
    while (left) {   // warp-uniform
        const int leader = __ffs(left) - 1;
        const int gb = __shfl_sync(0xffffffffu, b, leader);
        const bool mine = b == gb;
        const long long sum = k7_warp_sum_i64(mine ? iv : 0ll);
        if (lane == leader && sum) atomicAdd(reinterpret_cast<unsigned long long*>(acc + 2 * gb + which), (unsigned long long)sum);
        left &= ~__ballot_sync(0xffffffffu, mine);
    }
}

// ---- the exchange over peer memory.  Thread p posts this rank's sums into rank p's mailbox (16-byte store, system
// fence, then the sequence word with release semantics); a wait polls the own mailbox for rank p's post of the same
// call.  Two slot sets alternate by call parity: a rank can only be one call ahead of a peer (it needs the peer's
// post to finish - or, post_only, to collect - a call), so a slot is never overwritten before it was read.  Every
// rank adds the W posts in rank order: identical bits everywhere.
__device__ __forceinline__ void k7_peer_post(const K7PeerDev& PR, unsigned int seq, int p, const float4 mine) {
    const int W = PR.world;
    float* dst = PR.mailbox[p] + (size_t)((seq & 1u) * W + PR.rank) * 8;
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "f"(mine.x), "f"(mine.y), "f"(mine.z),
                 "f"(mine.w)
                 : "memory");
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(dst + 4), "r"(seq) : "memory");
}
__device__ __forceinline__ float4 k7_peer_wait(const K7PeerDev& PR, unsigned int seq, int p) {
    const float* src = PR.mailbox[PR.rank] + (size_t)((seq & 1u) * PR.world + p) * 8;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned int got = 0u;
    bool ok = false;
    for (;;) {
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(got) : "l"(src + 4) : "memory");
        if (got == seq) {
            ok = true;
            break;
        }
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > PR.timeout_ns) break;   // a missing peer must not hang the GPU: NaN sums instead
        __nanosleep(64);
    }
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                 : "l"(src)
                 : "memory");
    const float nanv = __int_as_float(0x7fc00000);
    return ok ? v : make_float4(nanv, nanv, nanv, nanv);
}
__device__ __forceinline__ void k7_peer_sum(const float4* got, int W, float* __restrict__ sums) {
    float4 a = got[0];
    for (int p = 1; p < W; ++p) {
        a.x += got[p].x;
        a.y += got[p].y;
        a.z += got[p].z;
        a.w += got[p].w;
    }
    sums[0] = a.x;
    sums[1] = a.y;
    sums[2] = a.z;
    sums[3] = a.w;
}

template <bool GRAD, bool FROM_SPEC, bool LEVELS>
__global__ void __launch_bounds__(K7_ROWS_THREADS)
k7_loss_serve_kernel(const __grid_constant__ K7AnchorSpec spec, const __grid_constant__ K7LevelTable LT,
                     const float* __restrict__ anchors, int box_order,
                     const float* __restrict__ cls, const float* __restrict__ reg, const float* __restrict__ annotations,
                     const __grid_constant__ K7RowQueues Q, const int32_t* __restrict__ num_pos,
                     const __grid_constant__ K7LossDev P, const __grid_constant__ K7PeerDev PR,
                     const float* __restrict__ pipe_partials, uint32_t pipe_slots,
                     long long* __restrict__ acc /* [B][2] */, unsigned int* __restrict__ counter,
                     float* __restrict__ per_image, float* __restrict__ sums, float* __restrict__ dcls,
                     float4* __restrict__ dreg) {
    constexpr int NW = K7_ROWS_THREADS / 32;
    __shared__ double s_red[2 * NW];
    __shared__ bool s_last;
    const int A = P.A, C = P.C;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float oma = P.one_m_alpha;
    const int img = blockIdx.y;   // grid (blocks per image, B)
    k7_pdl_trigger();
    k7_pdl_wait();      // the stream kernel's gradients and partial sums are complete
    const unsigned int npos = __ldcg(Q.count + (size_t)img * K7_BEST_CNT_STRIDE);

    // ---- positive rows of image `img`: a thread each (whole warps stay in the loop: the fold below is warp-wide)
    for (unsigned int i0 = blockIdx.x * K7_ROWS_THREADS; i0 < npos; i0 += gridDim.x * K7_ROWS_THREADS) {
        const unsigned int i = i0 + threadIdx.x;
        int b = -1;
        float cfix = 0.0f, rsum = 0.0f;
        if (i < npos) {
            const uint2 ent = __ldcg(Q.q + (size_t)img * A + i);
            const int a = (int)ent.x;
            b = img;
            const size_t R = (size_t)b * A + a;
            const float* arow = annotations + ((size_t)b * P.M + ent.y) * 5;
            const float2 sc = k7_loss_scales(__ldg(num_pos + b), P.mean_div);
            rsum = k7_row_smooth_l1<GRAD, FROM_SPEC>(spec, anchors, box_order, P, a, R, arow, reg, sc.y, dreg);
            const int label = (int)__ldg(arow + 4);   // .long() truncation, losses.py:97
            if (label >= 0 && label < C) {
                size_t e = R * (size_t)C + (size_t)label;
                const float* src = cls;
                float* dst = dcls;
                if (LEVELS) {   // NCHW per level: ((b*K*C + k*C + label) * S + cell)
                    int l = 0;
#pragma unroll
                    for (int i2 = 1; i2 < K7_MAX_LEVELS; ++i2)
                        if (i2 < spec.num_levels && a >= spec.offset[i2]) l = i2;
                    const int r = a - spec.offset[l], K = spec.cell_anchors;
                    const int cell = r / K, k = r - cell * K;
                    e = ((size_t)b * K * C + (size_t)k * C + (size_t)label) * (size_t)LT.S[l] + (size_t)cell;
                    src = LT.cls[l];
                    dst = LT.dcls[l];
                }
                const float zl = __ldg(src + e);
                const float xl = P.from_logits ? k7_sigmoid_precise(zl) : zl;
                const float p = k7_clamp_p(xl);
                cfix = k7_focal_pos(p, P.alpha) - oma * k7_neg_exact(xl, nullptr);
                if (GRAD) {
                    float g = (p == xl) ? k7_focal_pos_grad(p, P.alpha) * sc.x : 0.0f;
                    if (P.from_logits) g *= xl * (1.0f - xl);   // d/dz through the fused sigmoid
                    dst[e] = g;
                }
            }
        }
        k7_acc_add(acc, b, 0, (double)cfix, lane);
        k7_acc_add(acc, b, 1, (double)rsum, lane);
    }

    // ---- the last block to finish reduces
    __threadfence();   // this thread's accumulator updates are performed before the block's tick
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(counter, 1u) == gridDim.x * gridDim.y - 1u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // per-image losses (losses.py:118, 161-166): everything - the stream's share too - is in the two accumulators
    for (int b = threadIdx.x; b < P.B; b += K7_ROWS_THREADS) {
        const bool any_valid = Q.img_valid[b] != 0;
        const int n = num_pos[b];
        const double fc = (double)__ldcg(acc + 2 * b) * (1.0 / K7_FIX_ONE);
        const double fr = (double)__ldcg(acc + 2 * b + 1) * (1.0 / K7_FIX_ONE);
        per_image[2 * b + 0] = any_valid ? (float)fc / fmaxf((float)n, 1.0f) : 0.0f;
        per_image[2 * b + 1] = (any_valid && n > 0) ? (float)(fr / (4.0 * (double)n)) : 0.0f;
    }
    __syncthreads();

    // ---- batch sums in image order
    double c = 0.0, r = 0.0, nn = 0.0, zero = 0.0;
    for (int i = threadIdx.x; i < P.B; i += K7_ROWS_THREADS) {
        c += (double)per_image[2 * i];
        r += (double)per_image[2 * i + 1];
        nn += (double)num_pos[i];
    }
    k7_block_sum2_f64(c, r, s_red);
    k7_block_sum2_f64(nn, zero, s_red);
    __shared__ float4 s_mine, s_got[K7_MAX_PEERS];
    if (threadIdx.x == 0) {
        s_mine = make_float4((float)c, (float)r, (float)nn, (float)P.B);
        if (PR.world <= 1) {
            sums[0] = s_mine.x;
            sums[1] = s_mine.y;
            sums[2] = s_mine.z;
            sums[3] = s_mine.w;
        }
    }
    if (PR.world <= 1) return;

    // ---- the path's only exchange, fused: thread p posts this rank's sums into rank p's mailbox
    // (16-byte store, system fence, then the sequence word with release semantics) and waits for rank
    // p's post in its own mailbox.  Two slot sets alternate by call parity: a rank can only be one
    // call ahead of a peer (it needs the peer's post to finish a call), so a slot is never overwritten
    // before it was read.  Every rank adds the W posts in rank order: identical bits everywhere.
    // The call number: a kernel argument, or - so that a captured CUDA graph can be replayed - a device counter
    // this block advances (every rank makes the same calls, so the counters agree).
    __shared__ unsigned int s_seq;
    if (threadIdx.x == 0) s_seq = PR.seq_counter ? atomicAdd(PR.seq_counter, 1u) + 1u : PR.seq;
    __syncthreads();
    const unsigned int seq = s_seq;
    const int W = PR.world;
    if ((int)threadIdx.x < W) k7_peer_post(PR, seq, threadIdx.x, s_mine);
    if (PR.post_only) {   // nobody waits here: the rank's own sums for now, the global ones through k7_peer_collect
        if (threadIdx.x == 0) {
            sums[0] = s_mine.x;
            sums[1] = s_mine.y;
            sums[2] = s_mine.z;
            sums[3] = s_mine.w;
        }
        return;
    }
    if (threadIdx.x == 0 && PR.seq_counter) atomicAdd(PR.seq_counter + 1, 1u);   // this call collects itself
    if ((int)threadIdx.x < W) s_got[threadIdx.x] = k7_peer_wait(PR, seq, threadIdx.x);
    __syncthreads();
    if (threadIdx.x == 0) k7_peer_sum(s_got, W, sums);
}

// The second half of a post_only exchange (k7_peer_collect): W threads wait for the W posts of the next
// uncollected call, thread 0 adds them in rank order.
__global__ void k7_peer_collect_kernel(const __grid_constant__ K7PeerDev PR, float* __restrict__ sums) {
    __shared__ float4 s_got[K7_MAX_PEERS];
    __shared__ unsigned int s_seq;
    if (threadIdx.x == 0) s_seq = atomicAdd(PR.seq_counter + 1, 1u) + 1u;
    __syncthreads();
    const int W = PR.world;
    if ((int)threadIdx.x < W) s_got[threadIdx.x] = k7_peer_wait(PR, s_seq, threadIdx.x);
    __syncthreads();
    if (threadIdx.x == 0) k7_peer_sum(s_got, W, sums);
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

// several tensors in ONE launch (the two gradients of a step, or the per-level gradients): blockIdx.y = tensor
struct K7ScaleMulti {
    float* g[K7_SCALE_MAX];
    size_t n[K7_SCALE_MAX];
    const float* scale[K7_SCALE_MAX];
};
__global__ void k7_scale_multi_kernel(const __grid_constant__ K7ScaleMulti P) {
    const int t = blockIdx.y;
    const float s = __ldg(P.scale[t]);
    if (s == 1.0f) return;
    float* g = P.g[t];
    const size_t n = P.n[t], n4 = n >> 2;
    float4* g4 = reinterpret_cast<float4*>(g);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 v = g4[i];
        v.x *= s; v.y *= s; v.z *= s; v.w *= s;
        g4[i] = v;
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) g[(n4 << 2) + threadIdx.x] *= s;
}

extern "C" int k7_scale_inplace_multi(float* const* grads, const size_t* n, const float* const* scales, int count,
                                      void* stream) {
    K7_REQUIRE(grads && n && scales, "k7_scale_inplace_multi: null argument");
    K7_REQUIRE(count >= 0 && count <= K7_SCALE_MAX, "k7_scale_inplace_multi: count %d outside [0, %d]", count, K7_SCALE_MAX);
    K7ScaleMulti P;
    memset(&P, 0, sizeof(P));
    int m = 0;
    for (int i = 0; i < count; ++i) {
        if (n[i] == 0) continue;
        K7_REQUIRE(grads[i] && scales[i], "k7_scale_inplace_multi: null tensor %d", i);
        K7_REQUIRE((reinterpret_cast<uintptr_t>(grads[i]) & 15) == 0, "k7_scale_inplace_multi: grad %d must be 16-byte aligned", i);
        P.g[m] = grads[i];
        P.n[m] = n[i];
        P.scale[m] = scales[i];
        ++m;
    }
    if (m == 0) return K7_OK;
    k7_scale_multi_kernel<<<dim3(k7_sm_count() * 4, m), 256, 0, static_cast<cudaStream_t>(stream)>>>(P);
    K7_LAUNCH_CHECK();
    return K7_OK;
}

// ------------------------------------------------------------------ host side
#define K7_LOSS_MAX_GRID 256   // upper bound on CTAs assumed by the workspace query

// ------------------------------------------------------------------ workspace + launch plan
// One plan serves both entry points: k7_loss streams ONE tensor [B,A,C]; k7_loss_levels streams
// one NCHW tensor per pyramid level (see K7LevelTable).
struct K7LossPlan {
    int nseg;                               // streams of the one launch: 1, or the non-empty pyramid levels
    K7PipeGeom G;
    int level[K7_MAX_LEVELS];               // pyramid level of each stream (levels mode)
    unsigned long long AC[K7_MAX_LEVELS];   // elements per image of each stream
    uint32_t tile0[K7_MAX_LEVELS];
    size_t off_pipe;                        // per-warp slots of the launch
    size_t off_counter, off_acc, off_valid, off_poscnt, off_bits, off_cta, off_queue, total;
    size_t zero_bytes;                      // counters .. anchor bitmap: cleared at the start of every call
    size_t off_lbits[K7_MAX_LEVELS];        // per-level streams: the ignore bits expanded to one per element
    uint32_t cta_stride;
};

static int k7_pipe_geometry(bool grad, int B, int C, K7PipeGeom* g);
static int k7_env_int(const char* name, int dflt);

// levels == false: shape (B, A, C).  levels == true: one stream per non-empty level of `spec`.
static int k7_loss_plan(const K7AnchorSpec* spec, bool levels, int B, int A, int C, bool grad, bool exact_grid,
                        K7LossPlan* p, int spare_sms = 0) {
    p->nseg = 0;
    size_t o = 0;
    p->off_counter = o;   // [0] the serve kernel's block counter, [4], [5] the lengths of the two row queues
    o += 256;
    p->off_acc = o;       // [B][2] fixed-point row accumulators
    o += k7_align_up(sizeof(long long) * 2 * (size_t)B, 256);
    p->off_valid = o;     // [B] the image has an annotation row
    o += k7_align_up((size_t)B, 256);
    p->off_poscnt = o;    // [B] positives per image, K7_BEST_CNT_STRIDE words apart (the row queues' lengths)
    o += k7_align_up(sizeof(unsigned int) * K7_BEST_CNT_STRIDE * (size_t)B, 256);
    p->off_bits = o;      // [B][K7_IGNORE_WORDS(A)] one bit per anchor: the row is ignored (losses.py:79-82)
    o += k7_align_up(sizeof(uint32_t) * (size_t)B * K7_IGNORE_WORDS(A), 256);
    p->zero_bytes = o - p->off_counter;
    int rc = k7_pipe_geometry(grad, B, levels ? 1 : C, &p->G);
    if (rc) return rc;
    unsigned long long tiles = 0;
    const int nl = levels ? spec->num_levels : 1;
    for (int l = 0; l < nl; ++l) {
        unsigned long long ac = (unsigned long long)A * (unsigned long long)C;
        if (levels) {
            ac = (unsigned long long)spec->cell_anchors * C * spec->fh[l] * spec->fw[l];
            if (ac == 0) continue;
            K7_REQUIRE(ac < (1ull << 31), "k7_loss_levels: level %d has %llu elements per image", l, ac);
        }
        const int i = p->nseg++;
        p->level[i] = l;
        p->AC[i] = ac;
        p->tile0[i] = (uint32_t)tiles;
        tiles += (((ac * (unsigned long long)B) & ~3ull) + p->G.tile_f - 1) / p->G.tile_f;
        K7_REQUIRE(tiles < (1ull << 31), "k7_loss: %llu tiles exceed the 32-bit tile index", tiles);
        p->off_lbits[i] = o;
        if (levels) o += k7_align_up(sizeof(uint32_t) * (size_t)B * K7_IGNORE_WORDS(ac), 256);
    }
    p->G.num_tiles = (uint32_t)tiles;
    int grid = exact_grid ? k7_sm_count() : K7_LOSS_MAX_GRID;   // the workspace query has no device: assume the largest grid
    if (exact_grid) {   // K7LossParams.spare_sms: leave some SMs to kernels of other streams
        const int spare = k7_env_int("K7_PIPE_SPARE_SMS", spare_sms);
        if (spare > 0 && grid - spare >= 16) grid -= spare;
    }
    if (grid > K7_LOSS_MAX_GRID) grid = K7_LOSS_MAX_GRID;
    if (exact_grid && tiles < (unsigned long long)grid) grid = (int)tiles;
    p->G.grid = grid < 1 ? 1 : grid;
    p->G.cta_off = 0;
    p->G.cta_stride = p->cta_stride = (uint32_t)p->G.grid;
    p->off_pipe = o;
    o += k7_align_up(sizeof(float) * (size_t)B * (size_t)p->G.grid * K7_PIPE_MAX_CW, 256);
    p->off_cta = o;       // [B][cta_stride] stream partials, one per (image, CTA)
    o += k7_align_up(sizeof(float) * (size_t)B * p->cta_stride, 256);
    p->off_queue = o;     // [B*A] row queue: positives from the front, ignored rows from the back
    o += k7_align_up(sizeof(uint2) * (size_t)B * (size_t)A, 256);
    p->total = o;
    return K7_OK;
}

extern "C" size_t k7_loss_workspace_bytes(int B, int A, int C) {
    if (B <= 0 || A <= 0 || C <= 0) return 0;
    K7LossPlan p;
    if (k7_loss_plan(nullptr, false, B, A, C, true, false, &p)) return 0;
    return p.total;
}

extern "C" size_t k7_loss_levels_workspace_bytes(const K7AnchorSpec* spec, int B, int C) {
    if (!spec || B <= 0 || C <= 0 || k7_check_spec(spec, spec->total)) return 0;
    K7LossPlan p;
    if (k7_loss_plan(spec, true, B, spec->total, C, true, false, &p)) return 0;
    return p.total;
}

static int k7_env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    if (!v || !*v) return dflt;
    return atoi(v);
}

static int k7_pipe_geometry(bool grad, int B, int C, K7PipeGeom* g) {
    // Tuned on B200 (tools/tune_loss.py, C2 and C5): the forward pass wants long slices (12 warps x
    // 1024 floats: fewer per-slice waits); with gradients the stream is a copy and wants the most
    // warps on short slices (16 x 512) and no more than four 32 KB stages in flight.
    int slice = k7_env_int("K7_SLICE", grad ? 512 : 1024);
    slice = slice >= 1024 ? 1024 : 512;
    int cw = k7_env_int("K7_WARPS", slice == 1024 ? 12 : K7_PIPE_MAX_CW);
    if (cw > K7_PIPE_MAX_CW) cw = K7_PIPE_MAX_CW;
    if (cw < 1) cw = 1;
    g->cw = cw;
    g->slice = (uint32_t)slice;
    g->tile_f = (uint32_t)cw * (uint32_t)slice;
    g->div_c = k7_fastdiv_make((uint32_t)C);
    // rows touching a tile: (tile_f-1)/C + 2 at most; a part starts up to 127 bits into its first 16 bytes
    const uint32_t nrows_max = (g->tile_f - 1) / (uint32_t)C + 2;
    g->side_words = (((nrows_max + 254u) >> 7) + 1u) << 2;
    g->stage_stride = (uint32_t)k7_align_up((size_t)g->tile_f * 4 + (size_t)(2 * g->side_words + 4) * 4, 128);
    const size_t sco_bytes = (grad && B <= K7_PIPE_SCO_MAX) ? k7_align_up(sizeof(float) * (size_t)B, 16) : 0;
    const size_t budget = 226 * 1024 - sco_bytes;
    int stages = (int)((budget - 64) / ((size_t)g->stage_stride + sizeof(K7TileDesc) + 16));
    const int want = k7_env_int("K7_STAGES", 4);
    if (stages > want) stages = want;
    if (stages > 16) stages = 16;
    K7_REQUIRE(stages >= (grad ? 3 : 2), "k7_loss: a tile of %u bytes leaves no room for the ring", g->stage_stride);
    g->stages = stages;
    size_t o = (size_t)stages * g->stage_stride;
    g->off_desc = (uint32_t)o;
    o += (size_t)stages * sizeof(K7TileDesc);
    o = k7_align_up(o, 16);
    g->off_bars = (uint32_t)o;
    o += (size_t)stages * 16;
    g->off_sco = sco_bytes ? (uint32_t)o : 0u;
    o += sco_bytes;
    g->smem_bytes = o;
    return K7_OK;
}

// blocks per image of the per-image row kernels: about `per_sm` blocks per SM in total
static unsigned k7_rows_gx(int B, int per_sm) {
    const int want = per_sm * k7_sm_count();
    int gx = (want + B - 1) / B;
    return (unsigned)(gx < 1 ? 1 : gx);
}

static int k7_loss_run(const K7AnchorSpec* spec, const float* anchors, int box_order, const float* cls,
                       const float* const* levels, const float* reg, const float* annotations, const int32_t* assign,
                       const uint64_t* best, int32_t* num_pos, int B, int A, int C, int M,
                       const K7LossParams* params, float* per_image, float* sums, float* dcls, float* const* dlevels,
                       float* dreg, void* workspace, size_t workspace_bytes, void* stream) {
    const bool lv = levels != nullptr;
    const bool grad = dreg != nullptr;
    K7_REQUIRE(reg && annotations && num_pos && params && per_image && sums && workspace, "k7_loss: null argument");
    K7_REQUIRE(M <= 65535, "k7_loss: M > 65535 annotation rows per image not supported");
    K7_REQUIRE((assign != nullptr) != (best != nullptr), "k7_loss: give either k7_match's assign words or k7_match_scatter's best array");
    K7_REQUIRE(B > 0 && A > 0 && C > 0 && M > 0, "k7_loss: bad shape B=%d A=%d C=%d M=%d", B, A, C, M);
    K7_REQUIRE(B <= 65535, "k7_loss: B > 65535 not supported");
    K7_REQUIRE(params->mean_divisor > 0, "k7_loss: mean_divisor must be positive");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(reg) & 15) == 0 && (reinterpret_cast<uintptr_t>(dreg) & 15) == 0,
               "k7_loss: reg and dreg must be 16-byte aligned");
    if (best) {
        K7_REQUIRE(params->match_generation >= 1u && params->match_generation <= K7_BEST_GEN_MAX,
                   "k7_loss: match_generation %u outside [1, %u]", params->match_generation, K7_BEST_GEN_MAX);
        K7_REQUIRE(params->pos_thr >= params->neg_thr && params->neg_thr > 0.0f, "k7_loss: bad thresholds for the best array");
        K7_REQUIRE((reinterpret_cast<uintptr_t>(best) & 7) == 0, "k7_loss: best must be 8-byte aligned");
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);

    K7LossPlan L;
    int rc = k7_loss_plan(spec, lv, B, A, C, grad, true, &L, params->spare_sms);
    if (rc) return rc;
    if (workspace_bytes < L.total) {
        k7_set_error("k7_loss: workspace %zu < %zu bytes", workspace_bytes, L.total);
        return K7_EWORKSPACE;
    }
    unsigned char* ws = static_cast<unsigned char*>(workspace);
    K7_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "k7_loss: workspace must be 256-byte aligned");
    unsigned int* counter = reinterpret_cast<unsigned int*>(ws + L.off_counter);
    long long* acc = reinterpret_cast<long long*>(ws + L.off_acc);
    float* cta_partials = reinterpret_cast<float*>(ws + L.off_cta);

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
    P.mean_div = (float)params->mean_divisor;
    P.yx_targets = params->yx_targets;
    P.dreg_prezeroed = params->dreg_prezeroed;
    P.from_logits = params->from_logits ? 1 : 0;
    P.B = B; P.A = A; P.C = C; P.M = M;

    // counters + accumulators are contiguous at the start of the workspace
    K7_CUDA(cudaMemsetAsync(ws + L.off_counter, 0, L.zero_bytes, st));   // counters, accumulators, img_valid, the anchor bitmap
    K7_CUDA(cudaMemsetAsync(num_pos, 0, sizeof(int32_t) * (size_t)B, st));

    K7AssignSrc SRC;
    memset(&SRC, 0, sizeof(SRC));
    SRC.words = assign;
    SRC.best = reinterpret_cast<const unsigned long long*>(best);
    SRC.gen = params->match_generation;
    SRC.neg_thr = params->neg_thr;
    SRC.pos_thr = params->pos_thr;
    SRC.words_out = params->assign_out;
    SRC.consume = params->consume_best ? 1 : 0;
    K7RowQueues RQ;
    RQ.q = reinterpret_cast<uint2*>(ws + L.off_queue);
    RQ.count = reinterpret_cast<unsigned int*>(ws + L.off_poscnt);
    RQ.cap = (unsigned long long)B * (unsigned long long)A;
    RQ.img_valid = ws + L.off_valid;
    RQ.bits = reinterpret_cast<uint32_t*>(ws + L.off_bits);
    RQ.W = (int)K7_IGNORE_WORDS(A);
    float4* dreg4 = reinterpret_cast<float4*>(dreg);
    if (best && !SRC.words_out) {
        // the list: k7_match_scatter left the anchors it touched behind its array (K7_BEST_LIST_OFFSET)
        const unsigned char* raw = reinterpret_cast<const unsigned char*>(best);
        const unsigned int* tcount = reinterpret_cast<const unsigned int*>(raw + k7_best_list_off(B, A));
        const unsigned int* tlist = reinterpret_cast<const unsigned int*>(raw + k7_best_list_off(B, A) + k7_best_list_hdr(B));
        if (grad && !P.dreg_prezeroed) K7_CUDA(cudaMemsetAsync(dreg, 0, sizeof(float4) * (size_t)B * (size_t)A, st));
        k7_loss_list_kernel<<<dim3(k7_rows_gx(B, 4), B), K7_ROWS_THREADS, 0, st>>>(annotations, SRC, P, RQ, num_pos, tcount, tlist);
        K7_LAUNCH_CHECK();
    } else {   // the scan: num_pos, the row queues, the zeros of d(reg)
        const long long chunks = (long long)B * ((A + K7_ROWS_CHUNK - 1) / K7_ROWS_CHUNK);
        long long blocks = chunks;
        const long long cap = 8ll * k7_sm_count();   // eight resident blocks of 256 threads per SM; the rest is looped
        if (blocks > cap) blocks = cap;
        const dim3 sgrid((unsigned)blocks);
        if (grad) {
            if (best) k7_loss_scan_kernel<true, true><<<sgrid, K7_ROWS_THREADS, 0, st>>>(annotations, SRC, P, RQ, num_pos, dreg4);
            else      k7_loss_scan_kernel<true, false><<<sgrid, K7_ROWS_THREADS, 0, st>>>(annotations, SRC, P, RQ, num_pos, dreg4);
        } else {
            if (best) k7_loss_scan_kernel<false, true><<<sgrid, K7_ROWS_THREADS, 0, st>>>(annotations, SRC, P, RQ, num_pos, dreg4);
            else      k7_loss_scan_kernel<false, false><<<sgrid, K7_ROWS_THREADS, 0, st>>>(annotations, SRC, P, RQ, num_pos, dreg4);
        }
        K7_LAUNCH_CHECK();
    }

    // programmatic dependent launch of the stream and serve kernels: measured (K7_LOSS_PDL=1) -1..-4 us on the k7_loss
    // call alone, +15..25 us on the eager FocalLoss forward+backward (the extended launch costs the host more than
    // the GPU gains) - off by default
    static const bool loss_pdl = [] { const char* e = getenv("K7_LOSS_PDL"); return e ? atoi(e) != 0 : false; }();
    K7LevelTable LT;
    memset(&LT, 0, sizeof(LT));
    K7PipeSegs SG;
    memset(&SG, 0, sizeof(SG));
    SG.n = L.nseg;
    K7ExpandTable ET;
    memset(&ET, 0, sizeof(ET));
    int eblocks = 0;
    for (int i = 0; i < L.nseg; ++i) {
        K7PipeSeg& Q = SG.s[i];
        Q.cls = cls;
        Q.dcls = dcls;
        Q.bits = RQ.bits;
        Q.AC = L.AC[i];
        Q.N = Q.AC * (unsigned long long)B;
        Q.N4 = Q.N & ~3ull;
        Q.A = (uint32_t)A;
        Q.tile0 = L.tile0[i];
        if (lv) {
            const int l = L.level[i];
            const int S = spec->fh[l] * spec->fw[l], K = spec->cell_anchors;
            Q.cls = levels[l];
            Q.dcls = grad ? dlevels[l] : nullptr;
            K7_REQUIRE(Q.cls && (reinterpret_cast<uintptr_t>(Q.cls) & 15) == 0 &&
                           (!grad || (Q.dcls && (reinterpret_cast<uintptr_t>(Q.dcls) & 15) == 0)),
                       "k7_loss_levels: level %d: null or unaligned tensor", l);
            LT.cls[l] = Q.cls;
            LT.dcls[l] = Q.dcls;
            LT.S[l] = S;
            Q.A = (uint32_t)Q.AC;   // C == 1 mode: a row per element
            uint32_t* lbits = reinterpret_cast<uint32_t*>(ws + L.off_lbits[i]);
            Q.bits = lbits;
            const long long threads = (S & 31) == 0 ? (long long)K * (S >> 5) : ((long long)K * C * S + 31) / 32;
            ET.lbits[i] = lbits;
            ET.level_off[i] = spec->offset[l];
            ET.S[i] = S;
            ET.Wl[i] = (int)K7_IGNORE_WORDS(Q.A);
            ET.blk_off[i] = eblocks;
            eblocks += (int)((threads + 255) / 256);
        }
        Q.W = (uint32_t)K7_IGNORE_WORDS(Q.A);
    }
    if (lv && eblocks > 0) {
        for (int i = L.nseg; i <= K7_MAX_LEVELS; ++i) ET.blk_off[i] = eblocks;
        dim3 grid(eblocks, B);
        k7_bits_expand_kernel<<<grid, 256, 0, st>>>(RQ.bits, RQ.W, ET, spec->cell_anchors, C);
        K7_LAUNCH_CHECK();
    }
    if (L.G.num_tiles > 0 || L.nseg > 0) {
        const K7PipeGeom& G = L.G;
        float* pipe_partials = reinterpret_cast<float*>(ws + L.off_pipe);
        K7LossDev Ps = P;
        if (lv) {
            Ps.A = 0;   // per stream: K7PipeSeg::A
            Ps.C = 1;
        }
        auto launch = [&](auto kern) -> int {
            K7_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G.smem_bytes));
            k7_pdl_site() = loss_pdl ? 1 : 0;   // site 0: plain launch
            K7_CUDA(k7_launch_pdl(kern, dim3(G.grid), dim3((G.cw + 1) * 32), G.smem_bytes, st, SG, (const int32_t*)num_pos,
                                  (const unsigned char*)RQ.img_valid, Ps, G, pipe_partials, acc));
            return K7_OK;
        };
        const bool lg = P.from_logits != 0;
        if (G.slice == 1024) {
            if (lg) rc = grad ? launch(k7_loss_pipe_kernel<true, 1024, true>) : launch(k7_loss_pipe_kernel<false, 1024, true>);
            else    rc = grad ? launch(k7_loss_pipe_kernel<true, 1024, false>) : launch(k7_loss_pipe_kernel<false, 1024, false>);
        } else {
            if (lg) rc = grad ? launch(k7_loss_pipe_kernel<true, 512, true>) : launch(k7_loss_pipe_kernel<false, 512, true>);
            else    rc = grad ? launch(k7_loss_pipe_kernel<true, 512, false>) : launch(k7_loss_pipe_kernel<false, 512, false>);
        }
        if (rc) return rc;
        K7_CUDA(cudaGetLastError());
    }

    K7PeerDev PR;
    memset(&PR, 0, sizeof(PR));
    PR.world = 1;
    if (params->peer_reduce && params->peer_reduce->world > 1) {
        const K7PeerReduce* q = params->peer_reduce;
        K7_REQUIRE(q->world <= K7_MAX_PEERS && q->rank >= 0 && q->rank < q->world && (q->seq >= 1u || q->seq_counter),
                   "k7_loss: bad peer_reduce (world=%d rank=%d seq=%u)", q->world, q->rank, q->seq);
        PR.world = q->world;
        PR.rank = q->rank;
        PR.seq = q->seq;
        PR.seq_counter = q->seq_counter;
        PR.post_only = q->post_only ? 1 : 0;
        K7_REQUIRE(!PR.post_only || (q->seq_counter && (reinterpret_cast<uintptr_t>(q->seq_counter) & 3) == 0),
                   "k7_loss: peer_reduce.post_only needs seq_counter (two device words: calls, collects)");
        PR.timeout_ns = (unsigned long long)(q->timeout_ms > 0 ? q->timeout_ms : 10000) * 1000000ull;
        for (int i = 0; i < q->world; ++i) {
            K7_REQUIRE(q->mailbox[i] && (reinterpret_cast<uintptr_t>(q->mailbox[i]) & 31) == 0,
                       "k7_loss: peer_reduce mailbox %d is null or not 32-byte aligned", i);
            PR.mailbox[i] = q->mailbox[i];
        }
    }
    static const K7AnchorSpec kEmpty = {};
    const K7AnchorSpec& sp = spec ? *spec : kEmpty;
    const dim3 rgrid(k7_rows_gx(B, 4), B);   // the queue lengths are only known on the device: grid-stride loops per image
    k7_pdl_site() = loss_pdl ? 2 : 0;
#define K7_SERVE_LAUNCH(GR, SP, LV)                                                                                          \
    K7_CUDA(k7_launch_pdl(k7_loss_serve_kernel<GR, SP, LV>, rgrid, dim3(K7_ROWS_THREADS), 0, st, sp, LT, anchors, box_order, \
                          cls, reg, annotations, RQ, (const int32_t*)num_pos, P, PR, (const float*)cta_partials,              \
                          L.cta_stride, acc, counter, per_image, sums, dcls, dreg4))
    if (lv) { if (grad) K7_SERVE_LAUNCH(true, true, true); else K7_SERVE_LAUNCH(false, true, true); }
    else if (grad) { if (spec) K7_SERVE_LAUNCH(true, true, false); else K7_SERVE_LAUNCH(true, false, false); }
    else           { if (spec) K7_SERVE_LAUNCH(false, true, false); else K7_SERVE_LAUNCH(false, false, false); }
#undef K7_SERVE_LAUNCH
    K7_CUDA(cudaGetLastError());
    return K7_OK;
}

extern "C" int k7_loss(const K7AnchorSpec* spec, const float* anchors, int box_order, const float* cls, const float* reg,
                       const float* annotations, const int32_t* assign, const uint64_t* best,
                       int32_t* num_pos, int B, int A, int C, int M, const K7LossParams* params, float* per_image,
                       float* sums, float* dcls, float* dreg, void* workspace, size_t workspace_bytes, void* stream) {
    K7_REQUIRE(cls, "k7_loss: null argument");
    K7_REQUIRE(spec || anchors, "k7_loss: need an anchor spec or an anchor tensor");
    K7_REQUIRE((dcls == nullptr) == (dreg == nullptr), "k7_loss: dcls and dreg must be given together");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(cls) & 15) == 0 && (reinterpret_cast<uintptr_t>(dcls) & 15) == 0,
               "k7_loss: cls and dcls must be 16-byte aligned");
    if (spec) {
        int rc = k7_check_spec(spec, A);
        if (rc) return rc;
        box_order = spec->box_order;
    } else {
        K7_REQUIRE((reinterpret_cast<uintptr_t>(anchors) & 15) == 0, "k7_loss: anchors must be 16-byte aligned");
        K7_REQUIRE(box_order == K7_BOX_XYXY || box_order == K7_BOX_YXYX, "k7_loss: bad box_order");
    }
    return k7_loss_run(spec, anchors, box_order, cls, nullptr, reg, annotations, assign, best, num_pos, B, A, C, M,
                       params, per_image, sums, dcls, nullptr, dreg, workspace, workspace_bytes, stream);
}

extern "C" int k7_loss_levels(const K7AnchorSpec* spec, const float* const* levels, const float* reg,
                              const float* annotations, const int32_t* assign, const uint64_t* best,
                              int32_t* num_pos, int B, int C, int M, const K7LossParams* params, float* per_image,
                              float* sums, float* const* dlevels, float* dreg, void* workspace, size_t workspace_bytes,
                              void* stream) {
    K7_REQUIRE(spec && levels, "k7_loss_levels: an anchor spec and the level tensors are required");
    K7_REQUIRE((dlevels == nullptr) == (dreg == nullptr), "k7_loss_levels: dlevels and dreg must be given together");
    int rc = k7_check_spec(spec, spec->total);
    if (rc) return rc;
    return k7_loss_run(spec, nullptr, spec->box_order, nullptr, levels, reg, annotations, assign, best, num_pos, B,
                       spec->total, C, M, params, per_image, sums, nullptr, dlevels, dreg, workspace, workspace_bytes,
                       stream);
}

extern "C" int k7_peer_collect(const K7PeerReduce* q, float* sums, void* stream) {
    K7_REQUIRE(q && sums, "k7_peer_collect: null argument");
    K7_REQUIRE(q->world >= 1 && q->world <= K7_MAX_PEERS && q->rank >= 0 && q->rank < q->world && q->seq_counter,
               "k7_peer_collect: bad peer_reduce (world=%d rank=%d) or no seq_counter", q->world, q->rank);
    K7PeerDev PR;
    memset(&PR, 0, sizeof(PR));
    PR.world = q->world;
    PR.rank = q->rank;
    PR.seq_counter = q->seq_counter;
    PR.timeout_ns = (unsigned long long)(q->timeout_ms > 0 ? q->timeout_ms : 10000) * 1000000ull;
    for (int i = 0; i < q->world; ++i) {
        K7_REQUIRE(q->mailbox[i] && (reinterpret_cast<uintptr_t>(q->mailbox[i]) & 31) == 0,
                   "k7_peer_collect: mailbox %d is null or not 32-byte aligned", i);
        PR.mailbox[i] = q->mailbox[i];
    }
    k7_peer_collect_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(PR, sums);
    K7_LAUNCH_CHECK();
    return K7_OK;
}
