// Inference side: decode, clip, score threshold, compaction, sort, NMS.
//
//   k7_decode / k7_clip     BBoxTransform / ClipBoxes          architectures/utils.py:26-71
//                                                              (yx variant utils_efficientdet.py:8-50)
//   k7_detect               eval tail of every detector        architectures/retinanet.py:297-318,
//                                                              batched form utils/utils.py:76-114
//   k7_nms                  torchvision.ops.nms call sites     retinanet.py:314, utils/utils.py:93
//
// k7_detect is a chain of kernels (details at each of them):
//   1. k7_score_pipe_kernel streams classifications (one flat array of B*A*C floats) through a
//      warp-specialised TMA pipeline; per row it takes the class maximum, and only for rows above
//      the score threshold (strict '>') finds the first arg-max class, generates the anchor,
//      decodes and clips the box and records the survivor by anchor id.  Regression deltas are
//      read for survivors only.  Bytes: A*C*4 read per image + 32*S written.
//   2. the survivors' keys (score descending | anchor ascending) are sorted in chunks on all SMs and
//      merged by ranking (k7_nms_chunksort_kernel, k7_nms_rank_kernel); the ranked boxes are binned on
//      uniform grids by area class (k7_nms_cellorder_kernel); a thread per box tests the boxes of the
//      neighbouring cells that rank before it (k7_nms_pairs_kernel) and records the overlaps as edges;
//      k7_nms_resolve_kernel (one CTA per image) resolves greedy NMS as a fixed point over the edges.
//      The result is identical to the sequential algorithm.  Images with more survivors than the
//      workspace's cap go through k7_fallback_image (one CTA, chunks of 1024 boxes) inside the resolve kernel.
#include <math.h>
#include <string.h>
#include <mutex>

#include "k7_common.cuh"
#include "k7_nms_v1.cuh"

// ------------------------------------------------------------------ stand-alone decode / clip
template <bool FROM_SPEC>
__global__ void __launch_bounds__(256)
k7_decode_kernel(const __grid_constant__ K7AnchorSpec spec, const float* __restrict__ anchors, int box_order,
                 size_t anchor_batch_stride, const float4* __restrict__ deltas, int A,
                 const __grid_constant__ K7DecodeParams P, float4* __restrict__ boxes) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (a >= A) return;
    float ax1, ay1, ax2, ay2;
    k7_anchor<FROM_SPEC>(spec, anchors + (FROM_SPEC ? 0 : (size_t)b * anchor_batch_stride), box_order, a, ax1, ay1, ax2, ay2);
    const float4 d = __ldg(deltas + (size_t)b * A + a);
    K7Box o = (P.variant == K7_DECODE_RETINA) ? k7_decode_retina(ax1, ay1, ax2, ay2, d.x, d.y, d.z, d.w, P.mean, P.std)
                                              : k7_decode_effdet(ax1, ay1, ax2, ay2, d.x, d.y, d.z, d.w);
    if (P.clip) o = k7_clip_box(o, P.height, P.width);
    boxes[(size_t)b * A + a] = make_float4(o.x1, o.y1, o.x2, o.y2);
}

extern "C" int k7_decode(const K7AnchorSpec* spec, const float* anchors, int box_order, size_t anchor_batch_stride,
                         const float* deltas, int B, int A, const K7DecodeParams* params, float* boxes, void* stream) {
    K7_REQUIRE(deltas && params && boxes, "k7_decode: null argument");
    K7_REQUIRE(spec || anchors, "k7_decode: need an anchor spec or an anchor tensor");
    K7_REQUIRE(B > 0 && A > 0 && B <= 65535, "k7_decode: bad shape B=%d A=%d", B, A);
    K7_REQUIRE(params->variant == K7_DECODE_RETINA || params->variant == K7_DECODE_EFFDET, "k7_decode: bad variant");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(deltas) & 15) == 0 && (reinterpret_cast<uintptr_t>(boxes) & 15) == 0,
               "k7_decode: deltas and boxes must be 16-byte aligned");
    if (spec) {
        int rc = k7_check_spec(spec, A);
        if (rc) return rc;
    } else {
        K7_REQUIRE((reinterpret_cast<uintptr_t>(anchors) & 15) == 0, "k7_decode: anchors must be 16-byte aligned");
        K7_REQUIRE(anchor_batch_stride == 0 || anchor_batch_stride == (size_t)A * 4, "k7_decode: bad anchor batch stride");
    }
    dim3 grid((A + 255) / 256, B);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    static const K7AnchorSpec kEmpty = {};
    if (spec)
        k7_decode_kernel<true><<<grid, 256, 0, st>>>(*spec, nullptr, spec->box_order, 0, reinterpret_cast<const float4*>(deltas),
                                                     A, *params, reinterpret_cast<float4*>(boxes));
    else
        k7_decode_kernel<false><<<grid, 256, 0, st>>>(kEmpty, anchors, box_order, anchor_batch_stride,
                                                      reinterpret_cast<const float4*>(deltas), A, *params,
                                                      reinterpret_cast<float4*>(boxes));
    K7_LAUNCH_CHECK();
    return K7_OK;
}

__global__ void __launch_bounds__(256) k7_clip_kernel(float4* __restrict__ boxes, size_t n, int height, int width) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 v = boxes[i];
    K7Box b = {v.x, v.y, v.z, v.w};
    b = k7_clip_box(b, height, width);
    boxes[i] = make_float4(b.x1, b.y1, b.x2, b.y2);
}

extern "C" int k7_clip(float* boxes, int B, int A, int height, int width, void* stream) {
    K7_REQUIRE(boxes && B > 0 && A > 0, "k7_clip: bad argument");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15) == 0, "k7_clip: boxes must be 16-byte aligned");
    const size_t n = (size_t)B * A;
    k7_clip_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<float4*>(boxes), n, height, width);
    K7_LAUNCH_CHECK();
    return K7_OK;
}

// ------------------------------------------------------------------ workspace of k7_detect
struct K7DetectWs {
    // ---- zeroed at the start of every call (contiguous)
    int32_t* surv_count;    // [B] survivors (counted by the score kernel)
    int32_t* nedge;         // [B] edges found by the pair kernel
    unsigned int* bounds;   // [B][4] bounds of the live boxes' centres (k7_f2ord)
    uint32_t* bits;         // [B][W32] survivor bitmap by anchor id (for the survivor rank)
    uint2* ctab;            // [B][K7_GRID_CELLS] cell table
    // ---- not initialised
    int32_t* kept_count;    // [B]
    int32_t* nsel;          // [B] boxes in play (-1: fallback image)
    int32_t* ccnt;          // [B][K7_NMS_G]
    int32_t* wprefix;       // [B][W32] exclusive popcount prefix of bits
    uint64_t* keys0;        // [B][A] survivor keys in arrival order
    uint64_t* keys1;        // [B][A] keys in rank order
    uint64_t* keysC;        // [B][cap]
    uint64_t* ckeyC;        // [B][cap]
    float4* box;            // [B][A] decoded+clipped box of each survivor, by anchor id
    int32_t* label;         // [B][A] arg-max class of each survivor, by anchor id
    uint64_t* kept_key;     // [B][A] sort key of each kept box, in keep order
    float4* kept_box;       // [B][A]
    float4* rbox;           // [B][cap] boxes in rank order
    float4* cbox;           // [B][cap] boxes in cell order
    int32_t* crank;         // [B][cap]
    int32_t* clabel;        // [B][cap]
    int32_t* pend;          // [B][cap]
    unsigned int* edges;    // [B][ecap]
    void* ginfo;            // [B] K7Grid
    size_t bytes, zero_bytes;
    int W32;
    int cap, ecap;          // survivors per image the multi-SM NMS path is sized for; edges per image
    bool use_v1;            // many images: the per-image chain of k7_nms_v1.cu (its areas follow)
    K7V1Ws v1;
};
static int k7_nms_cap(int B, int n_max);

static K7DetectWs k7_detect_ws(void* base, int B, int A, bool with_v1 = true) {
    K7DetectWs w;
    unsigned char* p = static_cast<unsigned char*>(base);
    size_t o = 0;
    const size_t BA = (size_t)B * A;
    w.W32 = (A + 31) / 32;
    auto take = [&](size_t bytes) {
        unsigned char* r = p ? p + o : nullptr;
        o += k7_align_up(bytes, 256);
        return r;
    };
    w.cap = k7_nms_cap(B, A);
    w.ecap = w.cap * 16;   // K7_EDGE_PER_BOX
    const size_t cap = (size_t)w.cap;
    w.surv_count = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * B));
    w.nedge = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * B));
    w.bounds = reinterpret_cast<unsigned int*>(take(sizeof(unsigned int) * 4 * (size_t)B));
    w.bits = reinterpret_cast<uint32_t*>(take(sizeof(uint32_t) * (size_t)B * w.W32));
    w.ctab = reinterpret_cast<uint2*>(take(sizeof(uint2) * (size_t)B * (15 * 32 * 32)));   // K7_GRID_CELLS
    w.zero_bytes = o;
    w.kept_count = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * B));
    w.nsel = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * B));
    w.ccnt = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * 8 * (size_t)B));            // K7_NMS_G
    w.wprefix = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * (size_t)B * w.W32));
    w.keys0 = reinterpret_cast<uint64_t*>(take(sizeof(uint64_t) * BA));
    w.keys1 = reinterpret_cast<uint64_t*>(take(sizeof(uint64_t) * BA));
    w.keysC = reinterpret_cast<uint64_t*>(take(sizeof(uint64_t) * (size_t)B * cap));
    w.ckeyC = reinterpret_cast<uint64_t*>(take(sizeof(uint64_t) * (size_t)B * cap));
    w.box = reinterpret_cast<float4*>(take(sizeof(float4) * BA));
    w.label = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * BA));
    w.kept_key = reinterpret_cast<uint64_t*>(take(sizeof(uint64_t) * BA));
    w.kept_box = reinterpret_cast<float4*>(take(sizeof(float4) * BA));
    w.rbox = reinterpret_cast<float4*>(take(sizeof(float4) * (size_t)B * cap));
    w.cbox = reinterpret_cast<float4*>(take(sizeof(float4) * (size_t)B * cap));
    w.crank = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * (size_t)B * cap));
    w.clabel = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * (size_t)B * cap));
    w.pend = reinterpret_cast<int32_t*>(take(sizeof(int32_t) * (size_t)B * cap));
    w.edges = reinterpret_cast<unsigned int*>(take(sizeof(unsigned int) * (size_t)B * w.ecap));
    w.ginfo = take(320 * (size_t)B);   // sizeof(K7Grid) <= 320
    // the areas of the per-image chain are always there: the chain is a per-call choice (K7DetectParams.nms_chain)
    static const int v1_min = [] { const char* e = getenv("K7_V1_MIN_IMAGES"); return e ? atoi(e) : -1; }();
    w.use_v1 = v1_min >= 0 ? B >= v1_min : (B >= K7_V1_MIN_IMAGES || A < K7_V1_SMALL_IMAGE);   // the automatic choice
    w.v1.bytes = 0;
    if (with_v1) {   // (the stand-alone k7_nms family always runs the multi-SM chain)
        w.v1 = k7v1::workspace(p ? p + o : nullptr, B, A);
        o += w.v1.bytes;
    }
    w.bytes = o;
    return w;
}

extern "C" size_t k7_detect_workspace_bytes(int B, int A, int C) {
    (void)C;
    if (B <= 0 || A <= 0) return 0;
    return k7_detect_ws(nullptr, B, A).bytes;
}

// ------------------------------------------------------------------ kernel 1: score + decode + compaction
// The same warp-specialised TMA pipeline as k7_loss_pipe_kernel: one CTA per SM, a producer warp
// and 16 consumer warps, 32 KB stages; classifications is ONE flat array of B*A*C floats cut into
// tiles that ignore row and image boundaries (any shape streams through 16-byte bulk copies).  The
// consumer is light: a row survives iff ANY of its elements exceeds the threshold, so the flat
// pass is LDS.128 + three max + one compare per float4.  Every element above the threshold is a
// REPORT: the warp takes the report's row from the tile in shared memory, reduces its maximum and
// FIRST arg-max, and the report that sits exactly on the arg-max element queues the survivor -
// every surviving row has exactly one such element, so there is no duplicate and no atomic slot
// allocation.  The queue (shared memory, per warp) is drained at the end, one survivor per lane:
// anchor, regression deltas, decode + clip, and the records by anchor id (bitmap bit, score,
// label, box) from which k7_sort_kernel later ranks the survivors.  Nothing with a long latency
// sits between a stage becoming full and its release.
//
// LOGITS (K7DetectParams.from_logits): the tensor holds pre-sigmoid logits.  The sigmoid is
// monotone, so reports are taken against logit(thr) lowered by a safety margin and rows are
// reduced in logit space; the drain then computes the score 1/(1+expf(-z)) of the row maximum
// exactly as torch.sigmoid does, re-applies the strict threshold on that float32 probability, and
// moves the label to an EARLIER class whose logit is lower but rounds to the same probability
// (what torch.max over the probabilities would return).
struct K7ScoreGeom {
    unsigned long long N, N4;   // B*A*C and its multiple-of-4 part (the last N % 4 floats are served from global memory)
    uint32_t AC, tile_f, num_tiles;
    K7FastDiv div_c;
    uint32_t stage_stride, off_desc, off_bars, off_queue;
    int stages, cw, grid;
    size_t smem_bytes;
};
struct K7ScoreDesc {
    uint32_t n, b0, ge0, split;   // floats | image of float 0 | its element index inside that image | floats that belong to image b0
};
#define K7_SPIPE_SLICE 512
#define K7_SPIPE_MAX_CW 16
#define K7_SPIPE_QCAP 96    // queued survivors per consumer warp

template <bool FROM_SPEC, bool LOGITS>
__global__ void __launch_bounds__((K7_SPIPE_MAX_CW + 1) * 32, 1)
k7_score_pipe_kernel(const __grid_constant__ K7AnchorSpec spec, const float* __restrict__ anchors, int box_order,
                     const float* __restrict__ cls, const float4* __restrict__ reg, int A, int C, float report_thr,
                     const __grid_constant__ K7DetectParams P, const __grid_constant__ K7ScoreGeom G,
                     uint32_t* __restrict__ bits, int W32, float4* __restrict__ box_ws, int32_t* __restrict__ label_ws,
                     int32_t* __restrict__ surv_count, uint64_t* __restrict__ keys0) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int NS = G.stages, CW = G.cw;
    constexpr int SL = K7_SPIPE_SLICE;
    k7_pdl_trigger();   // k7_sort_kernel (a dependent launch) may queue behind this grid's CTAs
    K7ScoreDesc* desc = reinterpret_cast<K7ScoreDesc*>(smem + G.off_desc);
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
    if (blockIdx.x >= G.num_tiles && blockIdx.x != 0) return;
    const uint32_t ntl = blockIdx.x < G.num_tiles ? (G.num_tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;

    if (warp == CW) {   // ---------------------------------------------------------------- producer
        if (lane != 0) return;
        auto load = [&](const uint32_t j) {
            const unsigned long long g = (unsigned long long)blockIdx.x + (unsigned long long)j * gridDim.x;
            const int s = (int)(j % (uint32_t)NS);
            const unsigned long long off = g * G.tile_f;
            const unsigned long long left = G.N4 - off;
            K7ScoreDesc d;
            d.n = left < G.tile_f ? (uint32_t)left : G.tile_f;
            d.b0 = (uint32_t)(off / G.AC);
            d.ge0 = (uint32_t)(off - (unsigned long long)d.b0 * G.AC);
            const uint32_t in_img = G.AC - d.ge0;
            d.split = in_img < d.n ? in_img : d.n;
            desc[s] = d;
            k7_mbar_expect_tx(full + s, d.n * 4u);
            k7_bulk_g2s(smem + (size_t)s * G.stage_stride, cls + off, d.n * 4u, full + s);
        };
        uint32_t jl = 0;
        for (; jl < ntl && jl < (uint32_t)NS; ++jl) load(jl);
        int s = 0;
        uint32_t phase = 0;
        for (uint32_t j = 0; j < ntl; ++j) {
            k7_mbar_wait(done + s, phase);
            if (jl < ntl) load(jl++);
            if (++s == NS) {
                s = 0;
                phase ^= 1u;
            }
        }
        return;
    }
    if (warp > CW) return;

    // ------------------------------------------------------------------------------------ consumers
    const float thr = report_thr;
    const uint32_t full_u32 = k7_smem_u32(full), done_u32 = k7_smem_u32(done);
    const int i_lo = warp * SL;
    // queue entry: x = image (bit 31: row still unresolved), y = anchor (or element index inside the
    // image when unresolved), z = label, w = score bits
    uint4* queue = reinterpret_cast<uint4*>(smem + G.off_queue) + warp * K7_SPIPE_QCAP;
    int qn = 0;

    // Deferred emission, one queue entry per lane: anchor, regression deltas, decode + clip, and the
    // by-anchor records.  Unresolved entries (rows straddling two tiles, the unaligned tail) first
    // take their row from global memory.
    auto drain = [&]() {
        __syncwarp();
        for (int base = 0; base < qn; base += 32) {   // warp-uniform
            bool ok = base + lane < qn;
            uint32_t b = 0u, a = 0u;
            int idx = 0;
            float m = 0.0f;
            if (ok) {
                const uint4 ent = queue[base + lane];
                b = ent.x & 0x7fffffffu;
                a = ent.y;
                idx = (int)(ent.z & 0x7fffffffu);
                bool tie_scan = (ent.z >> 31) != 0u;   // LOGITS: an earlier class may round to the same probability
                m = __uint_as_float(ent.w);
                if (ent.x >> 31) {
                    tie_scan = true;
                    a = k7_div(ent.y, G.div_c);
                    const int c = (int)(ent.y - a * (uint32_t)C);
                    const float* row = cls + (size_t)b * G.AC + (size_t)a * C;
                    m = -INFINITY;
                    idx = 0;
#pragma unroll 8
                    for (int q = 0; q < C; ++q) {
                        const float v = __ldg(row + q);
                        if (v > m) {
                            m = v;
                            idx = q;
                        }
                    }
                    ok = idx == c;
                }
                if (LOGITS && ok) {
                    const float p = k7_sigmoid_precise(m);
                    ok = p > P.score_thr;   // else: a report inside the safety margin of logit(thr)
                    // torch.max over the PROBABILITIES returns the first class with the maximal float32 value:
                    // an earlier class with a slightly lower logit may round to the same probability.  Such a
                    // logit lies within 2^-23/(1-p) of m; beyond z = 8 the plateau widens, so every earlier
                    // logit above 8 is checked.
                    if (ok && tie_scan) {   // rare: flagged when the row was reduced in shared memory
                        const float* row = cls + (size_t)b * G.AC + (size_t)a * C;
                        const float lo = fminf(m - 2e-3f, 8.0f);
                        for (int q = 0; q < idx; ++q) {
                            const float v = __ldg(row + q);
                            if (v >= lo && k7_sigmoid_precise(v) == p) {
                                idx = q;
                                break;
                            }
                        }
                    }
                    m = p;
                }
            }
            // one slot reservation per image and round: the lanes of an image add up (a same-address atomic per
            // lane serialises in the L2: 512 k of them cost the single-class workload 90 us)
            const unsigned peers = __match_any_sync(0xffffffffu, ok ? b : (0x80000000u | (uint32_t)lane));
            if (ok) {
                const int leader = __ffs(peers) - 1;
                int slot = 0;
                if (lane == leader) slot = atomicAdd(surv_count + b, __popc(peers));
                slot = __shfl_sync(peers, slot, leader) + __popc(peers & ((1u << lane) - 1u));
                float ax1, ay1, ax2, ay2;
                k7_anchor<FROM_SPEC>(spec, anchors, box_order, (int)a, ax1, ay1, ax2, ay2);
                const size_t oa = (size_t)b * A + a;
                const float4 dl = __ldg(reg + oa);
                K7Box o = (P.decode.variant == K7_DECODE_RETINA)
                              ? k7_decode_retina(ax1, ay1, ax2, ay2, dl.x, dl.y, dl.z, dl.w, P.decode.mean, P.decode.std)
                              : k7_decode_effdet(ax1, ay1, ax2, ay2, dl.x, dl.y, dl.z, dl.w);
                if (P.decode.clip) o = k7_clip_box(o, P.decode.height, P.decode.width);
                box_ws[oa] = make_float4(o.x1, o.y1, o.x2, o.y2);
                label_ws[oa] = idx;
                // the survivor's key joins the image's list (arrival order; the sort fixes the order)
                keys0[(size_t)b * A + slot] = ((uint64_t)k7_score_key_desc(m) << 32) | (uint32_t)a;
                atomicOr(bits + (size_t)b * W32 + (a >> 5), 1u << (a & 31));
            }
        }
        __syncwarp();
        qn = 0;
    };

    int s = -1;
    uint32_t phase = 1u;
    for (uint32_t j = 0; j < ntl; ++j) {
        if (++s == NS) s = 0;
        if (s == 0) phase ^= 1u;
        const float* tile = reinterpret_cast<const float*>(smem + (size_t)s * G.stage_stride);
        k7_mbar_wait_u32(full_u32 + 8u * s, phase);
        const uint4 d = *reinterpret_cast<const uint4*>(desc + s);   // n, b0, ge0, split
        const int n = (int)d.x;
        if (i_lo < n) {
            const int len = min(n - i_lo, SL), n4 = len >> 2;
            const float4* t4 = reinterpret_cast<const float4*>(tile + i_lo);
            uint32_t hits = 0u;   // bit k: float4 lane+32k holds an element above the threshold
            if (n4 == SL / 4) {
#pragma unroll
                for (int k = 0; k < SL / 128; ++k) {
                    const float4 v = t4[lane + 32 * k];
                    if (fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)) > thr) hits |= 1u << k;
                }
            } else {
                for (int k = 0; lane + 32 * k < n4; ++k) {
                    const float4 v = t4[lane + 32 * k];
                    if (fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)) > thr) hits |= 1u << k;
                }
            }
            if (C == 1 && __any_sync(0xffffffffu, hits != 0u)) {
                // single-class heads (KSevenDet thermal faces): a row is one element, every report IS a survivor
                // with label 0.  Lanes queue their own reports (ballot prefix) instead of the warp resolving
                // them one by one - that loop was 170 us for 42 MB at 256 dense images.
                for (int k = 0; k < SL / 128; ++k) {
                    const bool mine = (hits >> k) & 1u;
                    if (!__any_sync(0xffffffffu, mine)) continue;
                    float v[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
                    if (mine) {
                        const float4 t = t4[lane + 32 * k];
                        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
                    }
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const bool rep = v[e] > thr;
                        const unsigned mb = __ballot_sync(0xffffffffu, rep);
                        if (!mb) continue;
                        const int need = __popc(mb);
                        if (qn + need > K7_SPIPE_QCAP) drain();
                        if (rep) {
                            const int i = i_lo + 4 * (lane + 32 * k) + e;
                            uint32_t bimg = d.y, ge = d.z + (uint32_t)i;
                            if ((uint32_t)i >= d.w) {
                                const uint32_t r = (uint32_t)i - d.w, qi = r / G.AC;
                                bimg = d.y + 1u + qi;
                                ge = r - qi * G.AC;
                            }
                            queue[qn + __popc(mb & ((1u << lane) - 1u))] = make_uint4(bimg, ge, 0u, __float_as_uint(v[e]));
                        }
                        qn += need;
                    }
                }
            } else if (__any_sync(0xffffffffu, hits != 0u)) {   // ~1 slice in 10
                // Every element above the threshold is a report.  The warp resolves it at once from the
                // tile: class maximum and FIRST arg-max of its row (torch.max, retinanet.py:300,316);
                // the report that sits on the arg-max element queues the survivor.
                for (int k = 0; k < SL / 128; ++k) {
                    const bool mine = (hits >> k) & 1u;
                    if (!__any_sync(0xffffffffu, mine)) continue;
                    float v[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
                    if (mine) {
                        const float4 t = t4[lane + 32 * k];
                        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
                    }
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        unsigned mb = __ballot_sync(0xffffffffu, v[e] > thr);
                        while (mb) {
                            const int src = __ffs(mb) - 1;
                            mb &= mb - 1;
                            const int i = i_lo + 4 * (src + 32 * k) + e;      // tile-local element index
                            uint32_t bimg = d.y, ge = d.z + (uint32_t)i;      // image, element index inside it
                            if ((uint32_t)i >= d.w) {                         // the tile runs into the next image(s)
                                const uint32_t r = (uint32_t)i - d.w, qi = r / G.AC;
                                bimg = d.y + 1u + qi;
                                ge = r - qi * G.AC;
                            }
                            const uint32_t a = k7_div(ge, G.div_c);
                            const int c = (int)(ge - a * (uint32_t)C);
                            const int i0 = i - c;                             // tile-local index of the row's class 0
                            if (qn == K7_SPIPE_QCAP) drain();
                            if (i0 < 0 || i0 + C > n) {                       // row straddles two tiles: resolve later
                                if (lane == 0) queue[qn] = make_uint4(bimg | 0x80000000u, ge, 0u, 0u);
                                ++qn;
                                continue;
                            }
                            const float* row = tile + i0;
                            float m = -INFINITY;
                            int idx = 0x7fffffff;
                            for (int q = lane; q < C; q += 32) {
                                const float x = row[q];
                                if (x > m) {
                                    m = x;
                                    idx = q;
                                }
                            }
#pragma unroll
                            for (int o = 16; o > 0; o >>= 1) {
                                const float om = __shfl_xor_sync(0xffffffffu, m, o);
                                const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
                                if (om > m || (om == m && oi < idx)) {
                                    m = om;
                                    idx = oi;
                                }
                            }
                            if (idx != c) continue;   // warp-uniform
                            uint32_t flag = 0u;
                            if (LOGITS) {   // does an earlier class sit on the same float32-sigmoid plateau as the maximum?
                                const float lo = fminf(m - 2e-3f, 8.0f);
                                bool cnd = false;
                                for (int q = lane; q < idx; q += 32) cnd |= row[q] >= lo;
                                flag = __any_sync(0xffffffffu, cnd) ? 0x80000000u : 0u;
                            }
                            if (lane == 0) queue[qn] = make_uint4(bimg, a, (uint32_t)idx | flag, __float_as_uint(m));
                            ++qn;
                        }
                    }
                }
            }
        }
        __syncwarp();
        if (lane == 0) k7_mbar_arrive_u32(done_u32 + 8u * s);
    }
    // the last N % 4 floats cannot be bulk-copied: CTA 0, warp 0 turns them into unresolved reports
    if (blockIdx.x == 0 && warp == 0 && G.N4 < G.N) {
        for (unsigned long long e = G.N4; e < G.N; ++e) {   // warp-uniform
            if (!(__ldg(cls + e) > thr)) continue;
            const uint32_t bimg = (uint32_t)(e / G.AC);
            const uint32_t ge = (uint32_t)(e - (unsigned long long)bimg * G.AC);
            if (qn == K7_SPIPE_QCAP) drain();
            if (lane == 0) queue[qn] = make_uint4(bimg | 0x80000000u, ge, 0u, 0u);
            ++qn;
        }
    }
    drain();
}

static int k7_score_geometry(int B, int A, int C, K7ScoreGeom* g) {
    const unsigned long long ac = (unsigned long long)A * (unsigned long long)C;
    K7_REQUIRE(ac < (1ull << 31), "k7_detect: A*C=%llu per image is too large", ac);
    g->AC = (uint32_t)ac;
    g->N = ac * (unsigned long long)B;
    g->N4 = g->N & ~3ull;
    g->cw = K7_SPIPE_MAX_CW;
    g->tile_f = (uint32_t)g->cw * K7_SPIPE_SLICE;
    const unsigned long long nt = (g->N4 + g->tile_f - 1) / g->tile_f;
    K7_REQUIRE(nt < (1ull << 31), "k7_detect: %llu tiles exceed the 32-bit tile index", nt);
    g->num_tiles = (uint32_t)nt;
    g->div_c = k7_fastdiv_make((uint32_t)C);
    g->stage_stride = g->tile_f * 4;
    g->stages = 6;
    size_t o = (size_t)g->stages * g->stage_stride;
    g->off_desc = (uint32_t)o;
    o += (size_t)g->stages * sizeof(K7ScoreDesc);
    g->off_bars = (uint32_t)o;
    o += (size_t)g->stages * 16;
    g->off_queue = (uint32_t)o;
    o += (size_t)g->cw * K7_SPIPE_QCAP * sizeof(uint4);
    g->smem_bytes = o;
    int grid = k7_sm_count();
    if (nt < (unsigned long long)grid) grid = (int)nt;
    g->grid = grid < 1 ? 1 : grid;
    return K7_OK;
}

// ------------------------------------------------------------------ kernel 1': the same on per-level NCHW head outputs
// levels[l] = [B, K*C, S_l] (layout in k7_loss.cu): the C scores of anchor (s, k) are
// S_l floats apart, so one THREAD owns an anchor and walks its classes; consecutive threads own
// consecutive cells of one in-cell anchor plane, i.e. every load instruction of a warp reads 128
// contiguous bytes (512 with the vector walk).  Running maximum and FIRST arg-max (strict '>'), then
// exactly the survivor record of k7_score_pipe_kernel's drain.  Plain loads, eight in flight per
// thread: the kernel is a pure stream.
struct K7ScoreLevels {
    const float* cls[K7_MAX_LEVELS];
    int S[K7_MAX_LEVELS];
    int blk_off[K7_MAX_LEVELS + 1];    // levels with S % 4 == 0: first 256-thread tile (4 cells a thread) of each level
                                       // within an image; [K7_MAX_LEVELS] = tiles per image
    int sblk_off[K7_MAX_LEVELS + 1];   // the other levels: first block (a cell a thread); [K7_MAX_LEVELS] = grid.x
};

// survivor record of one anchor (shared by the vector and the scalar walk)
template <bool LOGITS>
__device__ __forceinline__ void k7_levels_emit(const K7AnchorSpec& spec, const K7DetectParams& P, const float* base, int S,
                                               float m, int idx, bool tie, int b, int a, int A, const float4* __restrict__ reg,
                                               uint32_t* __restrict__ bits, int W32, float4* __restrict__ box_ws,
                                               int32_t* __restrict__ label_ws, int32_t* __restrict__ surv_count, uint64_t* __restrict__ keys0) {
    if (LOGITS) {
        const float p = k7_sigmoid_precise(m);
        if (!(p > P.score_thr)) return;
        const float lo = fminf(m - 2e-3f, 8.0f);   // see k7_score_pipe_kernel: plateau of the float32 sigmoid
        for (int q = 0; tie && q < idx; ++q) {   // tie: an earlier class reached lo (tracked by the walk)
            const float v = __ldg(base + (size_t)q * S);
            if (v >= lo && k7_sigmoid_precise(v) == p) {
                idx = q;
                break;
            }
        }
        m = p;
    }
    float ax1, ay1, ax2, ay2;
    k7_anchor_from_spec(spec, a, ax1, ay1, ax2, ay2);
    const size_t oa = (size_t)b * A + a;
    const float4 dl = __ldg(reg + oa);
    K7Box o = (P.decode.variant == K7_DECODE_RETINA)
                  ? k7_decode_retina(ax1, ay1, ax2, ay2, dl.x, dl.y, dl.z, dl.w, P.decode.mean, P.decode.std)
                  : k7_decode_effdet(ax1, ay1, ax2, ay2, dl.x, dl.y, dl.z, dl.w);
    if (P.decode.clip) o = k7_clip_box(o, P.decode.height, P.decode.width);
    box_ws[oa] = make_float4(o.x1, o.y1, o.x2, o.y2);
    label_ws[oa] = idx;
    {   // the survivor's key joins the image's list (arrival order; the sort fixes the order)
        const int slot = atomicAdd(surv_count + b, 1);
        keys0[(size_t)b * A + slot] = ((uint64_t)k7_score_key_desc(m) << 32) | (uint32_t)a;
    }
    atomicOr(bits + (size_t)b * W32 + (a >> 5), 1u << (a & 31));
}

// The per-level walk.  A thread owns FOUR consecutive cells of one in-cell anchor plane (levels with
// S % 4 == 0; the others - odd little maps such as 7x7 - go through k7_score_levels_scalar_kernel)
// and walks the C class planes at stride S.  The kernel is persistent (2 CTAs per SM): CTA i takes the
// 256-thread tiles i, i + grid, ...; every thread copies its own 16 bytes per class into a shared-
// memory ring with cp.async and reads them back itself (no block barrier on the data), and the ring
// runs on across tile boundaries, so (G-1)*U 16-byte loads per thread stay in flight for the whole
// kernel.  (With register loads ptxas kept two in flight; with one short-lived block per tile the
// achieved occupancy was a third of the theoretical one.)
// Rows above the threshold are queued in shared memory and served by dense warps when the queue
// fills: anchor in float64, decode with two expf, the exact sigmoid ... ~800 instructions that would
// otherwise run with one or two active lanes in nearly every warp.
#define K7_LV_G 3
#define K7_LV_U 6
#define K7_LV_QCAP 2048
#define K7_LV_SMEM (K7_LV_G * K7_LV_U * 256 * 16)

// queue entry {k | s << 4, score bits, label | tie << 31, b | l << 24}
template <bool LOGITS>
__device__ __forceinline__ void k7_levels_flush(const K7AnchorSpec& spec, const K7ScoreLevels& LV, const K7DetectParams& P,
                                                const int4* s_q, int qn, int A, int C, const float4* __restrict__ reg,
                                                uint32_t* __restrict__ bits, int W32, float4* __restrict__ box_ws,
                                                int32_t* __restrict__ label_ws, int32_t* __restrict__ surv_count, uint64_t* __restrict__ keys0) {
    const int K = spec.cell_anchors;
    for (int i = threadIdx.x; i < qn; i += 256) {   // records are keyed by anchor id: the queue order does not matter
        const int4 ent = s_q[i];
        const int k = ent.x & 15, s = ent.x >> 4, b = ent.w & 0xffffff, l = ent.w >> 24;
        const int S = LV.S[l];
        const float* base = LV.cls[l] + ((size_t)b * K + k) * (size_t)C * (size_t)S + (size_t)s;
        k7_levels_emit<LOGITS>(spec, P, base, S, __int_as_float(ent.y), ent.z & 0x7fffffff, ent.z < 0, b,
                               spec.offset[l] + s * K + k, A, reg, bits, W32, box_ws, label_ws, surv_count, keys0);
    }
}

template <bool LOGITS>
__global__ void __launch_bounds__(256, 2)
k7_score_levels_kernel(const __grid_constant__ K7AnchorSpec spec, const __grid_constant__ K7ScoreLevels LV, int num_tiles,
                       const float4* __restrict__ reg, int A, int C, float report_thr,
                       const __grid_constant__ K7DetectParams P, uint32_t* __restrict__ bits, int W32,
                       float4* __restrict__ box_ws, int32_t* __restrict__ label_ws, int32_t* __restrict__ surv_count, uint64_t* __restrict__ keys0) {
    extern __shared__ float4 s_ring[];   // [K7_LV_G][K7_LV_U][256]
    __shared__ int4 s_q[K7_LV_QCAP];
    __shared__ int s_qn;
    const int K = spec.cell_anchors;
    const int tiles_per_image = LV.blk_off[K7_MAX_LEVELS];
    const int NG = (C + K7_LV_U - 1) / K7_LV_U;   // class groups per tile
    const int my_tiles = ((int)blockIdx.x < num_tiles) ? (num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    if (threadIdx.x == 0) s_qn = 0;
    __syncthreads();

    // tile v -> this thread's column of planes: image, level, in-cell anchor, first cell
    auto open_tile = [&](int v, int& b, int& l, int& k, int& s, int& S) -> const float* {
        b = v / tiles_per_image;
        const int x = v - b * tiles_per_image;
        l = 0;
#pragma unroll
        for (int i = 1; i < K7_MAX_LEVELS; ++i)
            if (x >= LV.blk_off[i]) l = i;
        S = LV.S[l];
        const int S4 = S >> 2;
        const int t = (x - LV.blk_off[l]) * 256 + (int)threadIdx.x;
        if (t >= K * S4) return nullptr;
        k = t / S4;
        s = (t - k * S4) << 2;
        return LV.cls[l] + ((size_t)b * K + k) * (size_t)C * (size_t)S + (size_t)s;
    };

    const uint32_t slot0 = (uint32_t)__cvta_generic_to_shared(s_ring) + threadIdx.x * 16u;
    // ---- issue cursor
    int it = 0, ig = 0, islot = 0;   // tile ordinal of this CTA, class group, ring group
    const float* ibase = nullptr;
    int iS = 0;
    auto issue = [&]() {
        if (it < my_tiles) {
            if (ig == 0) {
                int b_, l_, k_, s_;
                ibase = open_tile((int)blockIdx.x + it * (int)gridDim.x, b_, l_, k_, s_, iS);
            }
            if (ibase) {
                const int c0 = ig * K7_LV_U;
#pragma unroll
                for (int j = 0; j < K7_LV_U; ++j)
                    if (c0 + j < C)
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(slot0 + (uint32_t)((islot * K7_LV_U + j) * 4096)),
                                     "l"(ibase + (size_t)(c0 + j) * iS)
                                     : "memory");
            }
            if (++ig == NG) {
                ig = 0;
                ++it;
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");   // also when empty: the group count stays uniform
        islot = (islot + 1 == K7_LV_G) ? 0 : islot + 1;
    };
#pragma unroll
    for (int g = 0; g < K7_LV_G; ++g) issue();

    // ---- consume cursor
    int cslot = 0;
    for (int ct = 0; ct < my_tiles; ++ct) {
        int b, l, k = 0, s = 0, S;
        const bool live = open_tile((int)blockIdx.x + ct * (int)gridDim.x, b, l, k, s, S) != nullptr;
        float m[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
        float pm[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // largest value before the arg max
        int idx[4] = {0, 0, 0, 0};
        for (int c0 = 0; c0 < C; c0 += K7_LV_U) {
            asm volatile("cp.async.wait_group %0;" ::"n"(K7_LV_G - 1) : "memory");
            if (live) {
#pragma unroll
                for (int j = 0; j < K7_LV_U; ++j) {
                    if (c0 + j < C) {
                        const float4 v = s_ring[(cslot * K7_LV_U + j) * 256 + threadIdx.x];
                        if (v.x > m[0]) { pm[0] = m[0]; m[0] = v.x; idx[0] = c0 + j; }
                        if (v.y > m[1]) { pm[1] = m[1]; m[1] = v.y; idx[1] = c0 + j; }
                        if (v.z > m[2]) { pm[2] = m[2]; m[2] = v.z; idx[2] = c0 + j; }
                        if (v.w > m[3]) { pm[3] = m[3]; m[3] = v.w; idx[3] = c0 + j; }
                    }
                }
            }
            issue();   // refills the group just read
            cslot = (cslot + 1 == K7_LV_G) ? 0 : cslot + 1;
        }
        if (live) {
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (m[e] > report_thr) {
                    const int tie = (LOGITS && pm[e] >= fminf(m[e] - 2e-3f, 8.0f)) ? (int)0x80000000 : 0;
                    s_q[atomicAdd(&s_qn, 1)] = make_int4(k | ((s + e) << 4), __float_as_int(m[e]), idx[e] | tie, b | (l << 24));
                }
        }
        __syncthreads();
        const int qn = s_qn;
        if (qn > K7_LV_QCAP - 1024) {   // the next tile may add up to 1024
            k7_levels_flush<LOGITS>(spec, LV, P, s_q, qn, A, C, reg, bits, W32, box_ws, label_ws, surv_count, keys0);
            __syncthreads();
            if (threadIdx.x == 0) s_qn = 0;
            __syncthreads();
        }
    }
    k7_levels_flush<LOGITS>(spec, LV, P, s_q, s_qn, A, C, reg, bits, W32, box_ws, label_ws, surv_count, keys0);
}

// Levels whose S is not a multiple of four: one thread per anchor, register loads.  These maps are
// the small ones (7x7, 13x13 ...); grid (blocks, B).
template <bool LOGITS>
__global__ void __launch_bounds__(256)
k7_score_levels_scalar_kernel(const __grid_constant__ K7AnchorSpec spec, const __grid_constant__ K7ScoreLevels LV,
                              const float4* __restrict__ reg, int A, int C, float report_thr,
                              const __grid_constant__ K7DetectParams P, uint32_t* __restrict__ bits, int W32,
                              float4* __restrict__ box_ws, int32_t* __restrict__ label_ws, int32_t* __restrict__ surv_count, uint64_t* __restrict__ keys0) {
    const int b = blockIdx.y;
    int l = 0;
#pragma unroll
    for (int i = 1; i < K7_MAX_LEVELS; ++i)
        if ((int)blockIdx.x >= LV.sblk_off[i]) l = i;
    const int K = spec.cell_anchors, S = LV.S[l];
    const int t = ((int)blockIdx.x - LV.sblk_off[l]) * 256 + (int)threadIdx.x;
    if (t >= K * S) return;
    const int k = t / S, s = t - k * S;
    const float* base = LV.cls[l] + ((size_t)b * K + k) * (size_t)C * (size_t)S + (size_t)s;
    float m = -INFINITY, pm = -INFINITY;
    int idx = 0;
#pragma unroll 8
    for (int c = 0; c < C; ++c) {
        const float v = __ldg(base + (size_t)c * S);
        if (v > m) {
            pm = m;
            m = v;
            idx = c;
        }
    }
    if (m > report_thr)
        k7_levels_emit<LOGITS>(spec, P, base, S, m, idx, LOGITS && pm >= fminf(m - 2e-3f, 8.0f), b, spec.offset[l] + s * K + k, A,
                               reg, bits, W32, box_ws, label_ws, surv_count, keys0);
}

// ------------------------------------------------------------------ kernel 2: per-image sort + NMS
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
// Round-1 design: one CTA per image ranked and ordered the survivors (k7_sort_kernel), a bitmask kernel
// tested pairs of 64-box blocks that coarse summaries could not rule out, one CTA per image resolved the
// bit matrix.  ncu (profiles/r4a, r3c): at 8 images per GPU (a sharded EfficientDet-D7 batch) the two
// one-CTA-per-image kernels ran for 146 us on 8 of 148 SMs; at 256 small dense images (KSevenDet
// thermal) the block summaries ruled out next to nothing and the bitmask kernel did 2 M pair tests per
// image for ~60 k that matter.  Round 2: every stage runs on all SMs and work follows the boxes:
//   k7_nms_chunksort_kernel  grid (chunks, B): a CTA sorts 2048 of the image's keys (score descending,
//                            anchor ascending) in shared memory; bounds of the box centres (atomics);
//   k7_nms_rank_kernel       the rank of a key = its index in its chunk + its lower bound in every other
//                            sorted chunk (keys are unique) - a merge by ranking, no serial step; the box
//                            goes to its rank; its CELL on a uniform grid chosen by the binade of its area
//                            (cell size ~ the box size) becomes the next sort key;
//   k7_nms_cellorder_kernel  the same chunk sort + ranking on (cell, rank) keys: boxes of a cell end up
//                            contiguous and in rank order, a table holds each cell's [start, end);
//   k7_nms_pairs_kernel      a THREAD per box: IoU > thr needs overlap > thr * own extent in x and in y and
//                            areas within 1/thr of each other, so the partners of a box lie in the cells
//                            within max(1 - t, 1/(2t) - 1/2) own extents of its centre (half an extent at
//                            t = 0.5), in the grids of the neighbouring binades; candidates ranked before the box are tested with
//                            torchvision's arithmetic and every hit is an EDGE (box, suppressor);
//   k7_nms_resolve_kernel    one CTA per image: greedy NMS as a fixed point over the edges.
// Work is proportional to the pairs that can overlap; nothing of size n^2 exists any more (the round-1
// bit matrix took cap^2/8 bytes per image, 4 GB per batch).  The kept set is exactly torchvision.ops.nms's:
// an edge says "a higher ranked box overlaps me above thr", which is all the greedy rule looks at.
// Images with more survivors than the workspace's cap, more edges than its edge list, or a threshold
// below K7_GRID_MIN_THR are served by k7_fallback_image (chunked, one CTA) inside the resolve kernel.
#define K7_SORT_MAX 16384                       // largest cap
#define K7_NMS_CH 2048                          // keys per chunk CTA
#define K7_NMS_G (K7_SORT_MAX / K7_NMS_CH)      // chunks per image at most
#define K7_GRID_EMIN 6                          // classes: binade of the area, clamped to [2^6, 2^21)
#define K7_GRID_EMAX 20
#define K7_GRID_NCLS (K7_GRID_EMAX - K7_GRID_EMIN + 1)
#define K7_GRID_SIDE 32                         // a class's grid has at most 32 x 32 cells
#define K7_GRID_CELLS (K7_GRID_NCLS * K7_GRID_SIDE * K7_GRID_SIDE)
#define K7_GRID_MIN_THR 0.05f
#define K7_EDGE_PER_BOX 16                      // edge list: K7_EDGE_PER_BOX * cap entries per image

// upper bound on survivors per image the multi-SM path is sized for (multiple of K7_NMS_CH)
static int k7_nms_cap(int B, int n_max) {
    (void)B;
    long long cap = ((long long)n_max + K7_NMS_CH - 1) / K7_NMS_CH * K7_NMS_CH;
    if (cap > K7_SORT_MAX) cap = K7_SORT_MAX;
    return (int)cap;
}

// monotone map float -> uint32 for atomicMax; 0 is below every float (the memset value)
__device__ __forceinline__ unsigned int k7_f2ord(float f) {
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float k7_ord2f(unsigned int o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

// The grids of an image, derived by every kernel from the image's bounds in the same way.
struct K7Grid {
    float ox, oy;
    float inv_cw[K7_GRID_NCLS], inv_ch[K7_GRID_NCLS];
    int gx[K7_GRID_NCLS], gy[K7_GRID_NCLS], off[K7_GRID_NCLS];
};
// bounds: {max cx, max -cx, max cy, max -cy} as k7_f2ord values.  Called by one thread.
__device__ void k7_grid_setup(const unsigned int* __restrict__ bounds, K7Grid& g) {
    float x1 = 0.f, x0 = 0.f, y1 = 0.f, y0 = 0.f;
    if (bounds[0] && bounds[1] && bounds[2] && bounds[3]) {
        x1 = k7_ord2f(bounds[0]);
        x0 = -k7_ord2f(bounds[1]);
        y1 = k7_ord2f(bounds[2]);
        y0 = -k7_ord2f(bounds[3]);
    }
    const float ex = fmaxf(x1 - x0, 1e-3f), ey = fmaxf(y1 - y0, 1e-3f);
    g.ox = x0;
    g.oy = y0;
    int off = 0;
    for (int c = 0; c < K7_GRID_NCLS; ++c) {
        const float s = exp2f(0.5f * (float)(K7_GRID_EMIN + c + 1));   // ~ the extent of a box of this class
        const int gx = max(1, min(K7_GRID_SIDE, (int)ceilf(ex / s)));
        const int gy = max(1, min(K7_GRID_SIDE, (int)ceilf(ey / s)));
        g.gx[c] = gx;
        g.gy[c] = gy;
        g.inv_cw[c] = (float)gx / ex;
        g.inv_ch[c] = (float)gy / ey;
        g.off[c] = off;
        off += gx * gy;
    }
}
__device__ __forceinline__ int k7_grid_class(float area) {   // area > 0, finite
    const int e = (int)((__float_as_uint(area) >> 23) & 0xffu) - 127;
    return max(0, min(K7_GRID_NCLS - 1, e - K7_GRID_EMIN));
}
__device__ __forceinline__ bool k7_box_live(const float4 b, float& w, float& h, float& area) {
    w = K7_FSUB(b.z, b.x);
    h = K7_FSUB(b.w, b.y);
    area = K7_FMUL(w, h);
    // a box with an empty or non-finite extent never suppresses and is never suppressed (torchvision: overlap 0 or NaN)
    return (w > 0.0f) && (h > 0.0f) && (area > 0.0f) && (area < 3.0e38f) && (fabsf(b.x) < 1.0e18f) && (fabsf(b.y) < 1.0e18f);
}

// lower bound of `key` in sorted keys[0..n)
__device__ __forceinline__ int k7_lower_bound(const uint64_t* __restrict__ keys, int n, unsigned long long key) {
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if ((unsigned long long)__ldcg(keys + mid) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

struct K7NmsArgs {
    int A, cap, pre_topk, class_aware;
    float thr;
    int brute_n;                   // images with at most this many survivors: all pairs are tested (k7_nms_brute_kernel), no grid
    const int32_t* surv_count;     // [B] survivors
    uint64_t* keys0;               // [B][A] survivor keys, in arrival order
    uint64_t* keysC;               // [B][cap] chunk-sorted keys
    uint64_t* keys1;               // [B][A] keys in rank order
    const float4* box_ws;          // [B][A] boxes by anchor id
    const int32_t* label_ws;       // [B][A] classes by anchor id
    float4* rbox;                  // [B][cap] boxes in rank order
    unsigned int* bounds;          // [B][4]
    int32_t* nsel;                 // [B] boxes in play (after pre_topk), -1: fallback image
    int32_t* ccnt;                 // [B][K7_NMS_G] live boxes per chunk of the cell sort
    uint64_t* ckeyC;               // [B][cap] chunk-sorted (cell, rank) keys
    float4* cbox;                  // [B][cap] boxes in cell order
    int32_t* crank;                // [B][cap] rank of the box at each cell-order index
    int32_t* clabel;               // [B][cap] class of the box at each cell-order index (class-aware NMS)
    uint2* ctab;                   // [B][K7_GRID_CELLS] {cap - first index, last index + 1} of each cell, 0 = empty
    int32_t* pend;                 // [B][cap] suppressors of each box, by rank
    unsigned int* edges;           // [B][ecap] rank << 16 | suppressor rank
    int32_t* nedge;                // [B]
    int ecap;
    K7Grid* ginfo;                 // [B] the image's grids (written by the rank kernels, read by the pair kernel)
};

// ---- chunk sort of the score keys + bounds of the live boxes
// The chunk kernels are launched with a grid of a few CTAs per SM that loop over the (image, chunk) items: most
// items are empty (an image with 850 survivors has one chunk of the eight its workspace could hold), and 2048 CTAs
// of 1024 threads that only find that out cost 20 us per launch.
__device__ void k7_nms_chunksort_item(const K7NmsArgs& P, const int b, const int c) {
    __shared__ unsigned long long sk[K7_NMS_CH];
    __shared__ unsigned int s_b[4];
    const int tid = threadIdx.x, lane = tid & 31;
    const int n_all = P.surv_count[b];
    const bool small = n_all <= P.cap && P.thr >= K7_GRID_MIN_THR;
    if (!small || n_all <= K7_NMS_CH) return;   // one chunk: k7_nms_rank_kernel does everything in one CTA
    const int m = min(K7_NMS_CH, n_all - c * K7_NMS_CH);
    if (m <= 0) return;
    const size_t ob = (size_t)b * P.A;
    if (tid < 4) s_b[tid] = 0u;
    const int N = k7_bitonic_size(m);
    float mx = -3.0e38f, nx = -3.0e38f, my = -3.0e38f, ny = -3.0e38f;
    for (int i = tid; i < N; i += K7_NMS_THREADS) {
        unsigned long long key = ~0ull;
        if (i < m) {
            key = P.keys0[ob + (size_t)c * K7_NMS_CH + i];
            const float4 bx = __ldg(P.box_ws + ob + (uint32_t)(key & 0xffffffffull));
            float w, h, ar;
            if (k7_box_live(bx, w, h, ar)) {
                const float cx = 0.5f * (bx.x + bx.z), cy = 0.5f * (bx.y + bx.w);
                mx = fmaxf(mx, cx); nx = fmaxf(nx, -cx); my = fmaxf(my, cy); ny = fmaxf(ny, -cy);
            }
        }
        sk[i] = key;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        nx = fmaxf(nx, __shfl_xor_sync(0xffffffffu, nx, o));
        my = fmaxf(my, __shfl_xor_sync(0xffffffffu, my, o));
        ny = fmaxf(ny, __shfl_xor_sync(0xffffffffu, ny, o));
    }
    __syncthreads();
    if (lane == 0 && mx > -3.0e38f) {
        atomicMax(&s_b[0], k7_f2ord(mx)); atomicMax(&s_b[1], k7_f2ord(nx));
        atomicMax(&s_b[2], k7_f2ord(my)); atomicMax(&s_b[3], k7_f2ord(ny));
    }
    k7_bitonic_sort_smem(sk, N);   // begins and ends with the block in step
    if (tid < 4 && s_b[tid]) atomicMax(P.bounds + 4 * b + tid, s_b[tid]);
    for (int i = tid; i < m; i += K7_NMS_THREADS) P.keysC[(size_t)b * P.cap + (size_t)c * K7_NMS_CH + i] = sk[i];
}
__global__ void __launch_bounds__(K7_NMS_THREADS, 1)
k7_nms_chunksort_kernel(const __grid_constant__ K7NmsArgs P, int B) {
    k7_pdl_trigger();
    k7_pdl_wait();      // the score kernel's keys
    // boxes in play per image (after pre_nms_topk), -1: the image goes to the single-CTA path
    for (int b = blockIdx.x * K7_NMS_THREADS + threadIdx.x; b < B; b += gridDim.x * K7_NMS_THREADS) {
        const int n_all = P.surv_count[b];
        const bool small = n_all <= P.cap && P.thr >= K7_GRID_MIN_THR;
        P.nsel[b] = small ? ((P.pre_topk > 0 && n_all > P.pre_topk) ? P.pre_topk : n_all) : -1;
    }
    const int G = P.cap / K7_NMS_CH;
    if (G <= 1) return;
    for (int item = blockIdx.x; item < G * B; item += gridDim.x) {   // chunk-major: the busy items first, spread over the CTAs
        k7_nms_chunksort_item(P, item % B, item / B);
        __syncthreads();
    }
}

// ---- ranks, boxes in rank order, (cell, rank) keys sorted per chunk
// cell key of a live box (its grid cell << 32 | rank), ~0 for a box without partners
__device__ __forceinline__ unsigned long long k7_cell_key(const K7Grid& g, const float4 bx, int rank) {
    float w, h, ar;
    if (!k7_box_live(bx, w, h, ar)) return ~0ull;
    const int cl = k7_grid_class(ar);
    const float cx = 0.5f * (bx.x + bx.z), cy = 0.5f * (bx.y + bx.w);
    const int ix = max(0, min(g.gx[cl] - 1, (int)((cx - g.ox) * g.inv_cw[cl])));
    const int iy = max(0, min(g.gy[cl] - 1, (int)((cy - g.oy) * g.inv_ch[cl])));
    return ((unsigned long long)(unsigned int)(g.off[cl] + iy * g.gx[cl] + ix) << 32) | (unsigned int)rank;
}

// Images whose survivors fit ONE chunk (the usual case: a few hundred to 2048 boxes) are served by one CTA
// from the unsorted keys to the cell table - score sort, ranks, bounds, cell sort, cell order - with nothing
// but shared memory between the steps.  Larger images: this kernel ranks the chunk-sorted keys against the
// other chunks (all of the image's keys are staged in shared memory for the binary searches: in global memory
// a search is eleven dependent trips to the L2) and sorts its chunk's cell keys; k7_nms_cellorder_kernel merges.
template <bool MULTI>   // two launches: a CTA per one-chunk image, and looping CTAs with the staging buffer for the others
__device__ void k7_nms_rank_item(const K7NmsArgs& P, const int b, const int c, unsigned long long* s_all /* multi: all chunk-sorted keys */) {
    __shared__ unsigned long long sk[K7_NMS_CH];
    __shared__ K7Grid s_g;
    __shared__ unsigned int s_b[4];
    __shared__ int s_live;
    const int tid = threadIdx.x, lane = tid & 31;
    const int n = P.nsel[b];
    if (n < 0) return;
    const int n_all = P.surv_count[b];
    const bool single = n_all <= K7_NMS_CH;
    if (single == MULTI) {   // the other launch's image
        if (MULTI && c > 0 && tid == 0) P.ccnt[b * K7_NMS_G + c] = 0;
        return;
    }
    const int m = single ? n_all : min(K7_NMS_CH, n_all - c * K7_NMS_CH);
    if (m <= 0) {
        if (tid == 0) P.ccnt[b * K7_NMS_G + c] = 0;
        return;
    }
    const size_t ob = (size_t)b * P.A, oc = (size_t)b * P.cap;
    const int N = k7_bitonic_size(m);
    if (tid == 0) s_live = 0;
    if (tid < 4) s_b[tid] = 0u;
    constexpr int PER = K7_NMS_CH / K7_NMS_THREADS;   // keys per thread
    float4 bxs[PER];
    int ranks[PER];

    if (!MULTI) {
        // ---- score sort in place: rank = position
        for (int i = tid; i < N; i += K7_NMS_THREADS) sk[i] = (i < m) ? P.keys0[ob + i] : ~0ull;
        __syncthreads();
        k7_bitonic_sort_smem(sk, N);
        float mx = -3.0e38f, nx = -3.0e38f, my = -3.0e38f, ny = -3.0e38f;
#pragma unroll
        for (int u = 0; u < PER; ++u) {
            const int i = tid + u * K7_NMS_THREADS;
            ranks[u] = -1;
            if (i < m && i < n) {   // pre_nms_topk drops the rest
                const unsigned long long key = sk[i];
                const float4 bx = __ldg(P.box_ws + ob + (uint32_t)(key & 0xffffffffull));
                bxs[u] = bx;
                ranks[u] = i;
                P.keys1[ob + i] = key;
                P.rbox[oc + i] = bx;
                float w, h, ar;
                if (k7_box_live(bx, w, h, ar)) {
                    const float cx = 0.5f * (bx.x + bx.z), cy = 0.5f * (bx.y + bx.w);
                    mx = fmaxf(mx, cx); nx = fmaxf(nx, -cx); my = fmaxf(my, cy); ny = fmaxf(ny, -cy);
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            nx = fmaxf(nx, __shfl_xor_sync(0xffffffffu, nx, o));
            my = fmaxf(my, __shfl_xor_sync(0xffffffffu, my, o));
            ny = fmaxf(ny, __shfl_xor_sync(0xffffffffu, ny, o));
        }
        if (n_all <= P.brute_n) {   // few boxes: k7_nms_brute_kernel tests all pairs, no cells needed
            if (tid == 0) P.ccnt[b * K7_NMS_G] = 0;
            return;
        }
        if (lane == 0 && mx > -3.0e38f) {
            atomicMax(&s_b[0], k7_f2ord(mx)); atomicMax(&s_b[1], k7_f2ord(nx));
            atomicMax(&s_b[2], k7_f2ord(my)); atomicMax(&s_b[3], k7_f2ord(ny));
        }
        __syncthreads();   // also: every thread has taken its keys out of sk
        if (tid == 0) k7_grid_setup(s_b, s_g);
        __syncthreads();
        if (tid < (int)(sizeof(K7Grid) / 4)) reinterpret_cast<int*>(P.ginfo + b)[tid] = reinterpret_cast<const int*>(&s_g)[tid];
    } else {
        // ---- all chunk-sorted keys of the image -> shared memory; rank = index in the own chunk + lower bounds in the others
        const int nch = (n_all + K7_NMS_CH - 1) / K7_NMS_CH;
        for (int i = tid; i < n_all; i += K7_NMS_THREADS) s_all[i] = __ldcg(P.keysC + oc + i);
        if (tid == 0) k7_grid_setup(P.bounds + 4 * b, s_g);
        __syncthreads();
        if (c == 0 && tid < (int)(sizeof(K7Grid) / 4)) reinterpret_cast<int*>(P.ginfo + b)[tid] = reinterpret_cast<const int*>(&s_g)[tid];
#pragma unroll
        for (int u = 0; u < PER; ++u) {
            const int i = tid + u * K7_NMS_THREADS;
            ranks[u] = -1;
            if (i < m) {
                const unsigned long long key = s_all[c * K7_NMS_CH + i];
                int rank = i;
                for (int c2 = 0; c2 < nch; ++c2) {
                    if (c2 == c) continue;
                    const unsigned long long* ch = s_all + c2 * K7_NMS_CH;
                    int lo = 0, hi = min(K7_NMS_CH, n_all - c2 * K7_NMS_CH);
                    while (lo < hi) {
                        const int mid = (lo + hi) >> 1;
                        if (ch[mid] < key) lo = mid + 1; else hi = mid;
                    }
                    rank += lo;
                }
                if (rank < n) {   // pre_nms_topk drops the rest
                    const float4 bx = __ldg(P.box_ws + ob + (uint32_t)(key & 0xffffffffull));
                    bxs[u] = bx;
                    ranks[u] = rank;
                    P.keys1[ob + rank] = key;
                    P.rbox[oc + rank] = bx;
                }
            }
        }
    }

    // ---- cell keys of this CTA's boxes, sorted
    int live = 0;
    for (int i = tid; i < N; i += K7_NMS_THREADS) sk[i] = ~0ull;
    __syncthreads();
#pragma unroll
    for (int u = 0; u < PER; ++u) {
        const int i = tid + u * K7_NMS_THREADS;
        if (ranks[u] >= 0) {
            const unsigned long long ck = k7_cell_key(s_g, bxs[u], ranks[u]);
            sk[i] = ck;
            if (ck != ~0ull) ++live;
            else P.pend[oc + ranks[u]] = 0;   // in play, but it has no partner: kept
        }
    }
    if (live) atomicAdd(&s_live, live);
    __syncthreads();
    k7_bitonic_sort_smem(sk, N);
    const int nl = s_live;   // the live keys sort before the ~0 fillers
    if (tid == 0) P.ccnt[b * K7_NMS_G + c] = nl;
    if (MULTI) {
        for (int i = tid; i < nl; i += K7_NMS_THREADS) P.ckeyC[oc + (size_t)c * K7_NMS_CH + i] = sk[i];
        return;
    }
    // ---- one chunk: the sorted position IS the cell-order index
    uint2* tab = P.ctab + (size_t)b * K7_GRID_CELLS;
    for (int i = tid; i < nl; i += K7_NMS_THREADS) {
        const unsigned long long ck = sk[i];
        const int rank = (int)(unsigned int)(ck & 0xffffffffull);
        const unsigned int cell = (unsigned int)(ck >> 32);
        P.cbox[oc + i] = __ldcg(P.rbox + oc + rank);   // written by this CTA (L2)
        P.crank[oc + i] = rank;
        if (P.class_aware) P.clabel[oc + i] = __ldg(P.label_ws + ob + (uint32_t)(__ldcg(P.keys1 + ob + rank) & 0xffffffffull));
        if (i == 0 || (unsigned int)(sk[i - 1] >> 32) != cell) tab[cell].x = (unsigned int)(P.cap - i);
        if (i == nl - 1 || (unsigned int)(sk[i + 1] >> 32) != cell) tab[cell].y = (unsigned int)(i + 1);
    }
}
template <bool MULTI>
__global__ void __launch_bounds__(K7_NMS_THREADS, 1)
k7_nms_rank_kernel(const __grid_constant__ K7NmsArgs P, int B) {
    extern __shared__ __align__(16) unsigned char dyn_raw[];
    k7_pdl_trigger();
    k7_pdl_wait();
    const int G = MULTI ? P.cap / K7_NMS_CH : 1;
    for (int item = blockIdx.x; item < G * B; item += gridDim.x) {
        k7_nms_rank_item<MULTI>(P, item % B, item / B, reinterpret_cast<unsigned long long*>(dyn_raw));
        __syncthreads();
    }
}

// ---- boxes in cell order + the cell table
__global__ void __launch_bounds__(K7_NMS_THREADS, 1)
k7_nms_cellorder_kernel(const __grid_constant__ K7NmsArgs P, int B) {
    k7_pdl_trigger();
    k7_pdl_wait();
    const int tid = threadIdx.x;
    const int G0 = P.cap / K7_NMS_CH;
    for (int item = blockIdx.x; item < G0 * B; item += gridDim.x) {
    const int b = item % B, c = item / B;
    if (P.nsel[b] < 0 || P.surv_count[b] <= K7_NMS_CH) continue;   // one chunk: already in cell order
    const int m = P.ccnt[b * K7_NMS_G + c];
    if (m <= 0) continue;
    const size_t ob = (size_t)b * P.A, oc = (size_t)b * P.cap;
    const int G = G0;   // chunks of this workspace
    int cnt[K7_NMS_G];
#pragma unroll
    for (int c2 = 0; c2 < K7_NMS_G; ++c2) cnt[c2] = c2 < G ? P.ccnt[b * K7_NMS_G + c2] : 0;
    for (int i = tid; i < m; i += K7_NMS_THREADS) {
        const unsigned long long ck = __ldcg(P.ckeyC + oc + (size_t)c * K7_NMS_CH + i);
        int p = i;
#pragma unroll
        for (int c2 = 0; c2 < K7_NMS_G; ++c2)
            if (c2 != c && cnt[c2] > 0) p += k7_lower_bound(P.ckeyC + oc + (size_t)c2 * K7_NMS_CH, cnt[c2], ck);
        const int rank = (int)(unsigned int)(ck & 0xffffffffull);
        const unsigned int cell = (unsigned int)(ck >> 32);
        P.cbox[oc + p] = __ldcg(P.rbox + oc + rank);
        P.crank[oc + p] = rank;
        if (P.class_aware) P.clabel[oc + p] = __ldg(P.label_ws + ob + (uint32_t)(__ldcg(P.keys1 + ob + rank) & 0xffffffffull));
        uint2* t = P.ctab + (size_t)b * K7_GRID_CELLS + cell;
        atomicMax(&t->x, (unsigned int)(P.cap - p));   // first index of the cell, as cap - index
        atomicMax(&t->y, (unsigned int)(p + 1));       // last index + 1
    }
    }
}

// ---- pair tests.  The exact decision of torchvision.ops.nms is fl(inter / (area_i + area_j - inter)) > thr.
// Candidates first meet a cheap FILTER that no suppressed pair can fail:
//     w > 0, h > 0  and  inter > fl(c*area_i) + fl(c*area_j),   c = thr/(1+thr) * (1 - 1e-5).
// Proof: a suppressed pair has w, h > 0 (else inter is 0 or NaN and the quotient is not > thr >= 0),
// hence a positive union u, and fl(inter/u) > thr gives inter > thr*u*(1 - 2^-24), i.e.
// inter*(1 + thr') > thr'*(area_i + area_j)*(1 - 2^-23) with thr' = thr*(1 - 2^-23); the float32
// roundings of the two products and the sum are below 2^-22 relative, far inside the 1e-5 margin.
// Pairs that pass are re-tested exactly, division included.
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

// The decision with the division only where it matters.  Besides the necessary condition above there is a
// SUFFICIENT one: inter > K * (fl(c*area_i) + fl(c*area_j)) with K = 1.0000221 means
// inter > thr/(1+thr) * (1 + 1.19e-5) * (area_i + area_j), i.e. inter / u > thr * (1 + 1.19e-5) for the real union
// u = area_i + area_j - inter >= (area_i + area_j)/2 (inter <= min(area_i, area_j) holds for the rounded values
// too: rounding is monotone); the two roundings of u (<= 2^-21 of u) and the one of the quotient (2^-24) cannot
// bring it down to thr.  An overflowed inter is left to the exact test (inf / (inf - inf) is NaN: kept).  Dense
// clusters pass the filter by the thousand; only pairs within 2.2e-5 of the threshold reach the division.
__device__ __forceinline__ bool k7_nms_decide(const K7Box& a, float area_a, float ca, const float4 b, float cb, float thr) {
    const float w = K7_FSUB(fminf(a.x2, b.z), fmaxf(a.x1, b.x));
    const float h = K7_FSUB(fminf(a.y2, b.w), fmaxf(a.y1, b.y));
    if (!((w > 0.0f) && (h > 0.0f))) return false;
    const float inter = K7_FMUL(w, h), s = K7_FADD(ca, cb);
    if (!(inter > s)) return false;
    if (inter > K7_FMUL(s, 1.0000221f) && inter <= 3.0e38f) return true;
    return k7_nms_suppress_exact(a, area_a, b, k7_area(b.x, b.y, b.z, b.w), thr);
}

// Where the partners of box i can be.  IoU > t  =>  inter > t/(1+t) (A_i + A_j) and A_j > t A_i (inter <= A_j),
// so inter > t A_i; with inter = iw * ih and ih <= h_i:  iw > t w_i, and in the same way ih > t h_i, iw > t w_j,
// ih > t h_j.  With iw <= min(w_i, w_j) that gives w_j in (t w_i, w_i / t), and the centres satisfy
//     |cx_j - cx_i| <= (w_i + w_j)/2 - iw < (w_i + w_j)/2 - t max(w_i, w_j) <= k w_i,   k = max(1 - t, 1/(2t) - 1/2)
// (the middle term is largest at w_j = w_i for t >= 1/2 and at w_j = w_i / t below), the same in y, and the areas
// A_j in (t A_i, A_i / t).  At t = 0.5: centres within HALF an own extent, areas within a factor two.  All of it is
// evaluated with t lowered by 1e-4 relative and the radius widened by 1e-2 px, against float32 roundings of 1e-7.
// A CTA takes 256 consecutive boxes of the cell order.  Their partners lie, per class, in a band of grid rows - ONE
// contiguous run of the cell order: the runs (boxes, ranks, classes) and the rows' cell-table slices are staged in
// shared memory with coalesced loads, then every thread walks the cells of ITS OWN neighbourhood and tests the boxes
// that rank before its own - exact neighbourhoods (the work is the ~50 - 150 candidates a box really has), shared-
// memory latency.  History (profiles/r4g - r4i): a thread walking its cells in global memory spent 600 cycles of L2
// latency per candidate, twice (127 us for 54 k boxes); a warp testing the union of its lanes' neighbourhoods ran
// 8 x more tests than needed on small dense images (650 us at 256 x 2000 faces); and an atomicAdd on the image's
// edge counter per edge serialised in the L2 (10 k edges on 8 addresses cost more than all pair tests): hits now
// wait in a shared-memory buffer and leave with one reservation per CTA.
#define K7_PAIRS_THREADS 256
#define K7_PAIRS_SCAP 2048      // staged candidate boxes
#define K7_PAIRS_TCAP 1024      // staged cell-table entries
#define K7_PAIRS_EBUF 2048      // buffered edges, split between the groups
#define K7_PAIRS_GW 8           // lanes that share a box
#define K7_PAIRS_MAXCLS 6       // classes a CTA's boxes can look into: 3 - 4 binades for one box, one more across a boundary
struct K7PairsCls {
    int row0, nrows, gx, tab_off, run_lo, run_hi, box_off, pad;
};
#define K7_PAIRS_SMEM (K7_PAIRS_SCAP * (sizeof(float4) + 2 * sizeof(int)) + K7_PAIRS_TCAP * sizeof(uint2) + K7_PAIRS_EBUF * sizeof(unsigned int) + \
                       K7_PAIRS_THREADS * (sizeof(float4) + 2 * sizeof(int) + sizeof(uint2)))
__global__ void __launch_bounds__(K7_PAIRS_THREADS)
k7_nms_pairs_kernel(const __grid_constant__ K7NmsArgs P, float kq /* k of the bound above */, int B) {
    extern __shared__ __align__(16) unsigned char dyn_raw[];
    float4* s_cbox = reinterpret_cast<float4*>(dyn_raw);
    int* s_crank = reinterpret_cast<int*>(s_cbox + K7_PAIRS_SCAP);
    int* s_clab = s_crank + K7_PAIRS_SCAP;
    uint2* s_tab = reinterpret_cast<uint2*>(s_clab + K7_PAIRS_SCAP);
    unsigned int* s_edge = reinterpret_cast<unsigned int*>(s_tab + K7_PAIRS_TCAP);
    float4* s_mb = reinterpret_cast<float4*>(s_edge + K7_PAIRS_EBUF);       // the CTA's own boxes
    uint2* s_grun = reinterpret_cast<uint2*>(s_mb + K7_PAIRS_THREADS);      // run tables of the groups (32 entries per warp)
    int* s_mri = reinterpret_cast<int*>(s_grun + K7_PAIRS_THREADS);
    int* s_mlab = s_mri + K7_PAIRS_THREADS;
    __shared__ K7Grid s_g;
    __shared__ K7PairsCls s_cls[K7_PAIRS_MAXCLS];
    __shared__ int s_y0[K7_GRID_NCLS], s_y1[K7_GRID_NCLS];
    __shared__ int s_clo, s_chi, s_ok;
    const int tid = threadIdx.x;
    k7_pdl_trigger();
    k7_pdl_wait();
    const int per_img = P.cap / K7_PAIRS_THREADS;   // items (256 boxes of the cell order) per image
    const int BIG = 0x3fffffff;
    for (int item = blockIdx.x; item < per_img * B; item += gridDim.x) {   // block-uniform; block-major: the busy items first
        const int blk = item / B, b = item - blk * B;
        if (P.nsel[b] < 0) continue;
        int nc = 0;
        for (int c2 = 0; c2 < P.cap / K7_NMS_CH; ++c2) nc += P.ccnt[b * K7_NMS_G + c2];
        if (blk * K7_PAIRS_THREADS >= nc) continue;
        __syncthreads();   // the previous item is done with shared memory
        if (tid < (int)(sizeof(K7Grid) / 4)) reinterpret_cast<int*>(&s_g)[tid] = reinterpret_cast<const int*>(P.ginfo + b)[tid];
        if (tid < K7_GRID_NCLS) {
            s_y0[tid] = BIG;
            s_y1[tid] = -1;
        }
        if (tid == 0) {
            s_clo = BIG;
            s_chi = -1;
            s_ok = 1;
        }
        __syncthreads();
        const int p = blk * K7_PAIRS_THREADS + tid;
        const bool valid = p < nc;
        const size_t oc = (size_t)b * P.cap;
        const float4 mb = valid ? __ldcg(P.cbox + oc + p) : make_float4(0.f, 0.f, 0.f, 0.f);
        const int ri = valid ? __ldcg(P.crank + oc + p) : -1;
        const int mlabel = (valid && P.class_aware) ? __ldcg(P.clabel + oc + p) : 0;
        const float h = K7_FSUB(mb.w, mb.y), marea = K7_FMUL(K7_FSUB(mb.z, mb.x), h);
        const float thr = P.thr, tl = thr * (1.0f - 1e-4f);
        const float c_flt = thr / (1.0f + thr) * (1.0f - 1e-5f);
        const float cy = 0.5f * (mb.y + mb.w);
        const float ry = kq * h + 1e-2f;
        const int clo = valid ? k7_grid_class(fmaxf(marea * tl * (1.0f - 1e-5f), 1e-30f)) : BIG;
        const int chi = valid ? k7_grid_class(fminf(marea / tl * (1.0f + 1e-5f), 3.0e38f)) : -1;
        // ---- the CTA's bands: classes, and per class the rows its boxes look into
        if (valid) {
            atomicMin(&s_clo, clo);
            atomicMax(&s_chi, chi);
            for (int cl = clo; cl <= chi; ++cl) {
                const int gy = s_g.gy[cl];
                const float ich = s_g.inv_ch[cl];
                atomicMin(&s_y0[cl], max(0, min(gy - 1, (int)floorf((cy - ry - s_g.oy) * ich))));
                atomicMax(&s_y1[cl], max(0, min(gy - 1, (int)floorf((cy + ry - s_g.oy) * ich))));
            }
        }
        __syncthreads();
        const int CLO = s_clo, CHI = s_chi;
        const uint2* tab = P.ctab + (size_t)b * K7_GRID_CELLS;
        if (tid == 0) {   // layout of the staged table slices
            int toff = 0;
            if (CHI - CLO + 1 > K7_PAIRS_MAXCLS) s_ok = 0;
            for (int cl = CLO; cl <= CHI && cl - CLO < K7_PAIRS_MAXCLS; ++cl) {
                K7PairsCls& D = s_cls[cl - CLO];
                D.gx = s_g.gx[cl];
                D.row0 = s_y0[cl];
                D.nrows = max(0, s_y1[cl] - s_y0[cl] + 1);
                D.tab_off = toff;
                D.run_lo = BIG;
                D.run_hi = 0;
                toff += D.nrows * D.gx;
            }
            if (toff > K7_PAIRS_TCAP) s_ok = 0;
        }
        __syncthreads();
        const bool staged_tab = s_ok != 0;
        if (staged_tab) {   // the table slices, and from them each class's run of the cell order
            for (int cl = CLO; cl <= CHI; ++cl) {
                K7PairsCls& D = s_cls[cl - CLO];
                const int ne = D.nrows * D.gx;
                const uint2* src = tab + s_g.off[cl] + D.row0 * D.gx;
                int lo = BIG, hi = 0;
                for (int i = tid; i < ne; i += K7_PAIRS_THREADS) {
                    const uint2 se = __ldcg(src + i);
                    s_tab[D.tab_off + i] = se;
                    if (se.y) {
                        lo = min(lo, P.cap - (int)se.x);
                        hi = max(hi, (int)se.y);
                    }
                }
                if (hi) {
                    atomicMin(&D.run_lo, lo);
                    atomicMax(&D.run_hi, hi);
                }
            }
        }
        __syncthreads();
        if (tid == 0 && staged_tab) {   // layout of the staged runs
            int boff = 0;
            for (int cl = CLO; cl <= CHI; ++cl) {
                K7PairsCls& D = s_cls[cl - CLO];
                D.box_off = boff;
                boff += max(0, D.run_hi - D.run_lo);
            }
            if (boff > K7_PAIRS_SCAP) s_ok = 0;
        }
        __syncthreads();
        const bool staged = s_ok != 0;   // else: a pathological item, walked in global memory
        if (staged) {
            for (int cl = CLO; cl <= CHI; ++cl) {
                const K7PairsCls D = s_cls[cl - CLO];
                const int nb = D.run_hi - D.run_lo;
                for (int i = tid; i < nb; i += K7_PAIRS_THREADS) {
                    s_cbox[D.box_off + i] = __ldcg(P.cbox + oc + D.run_lo + i);
                    s_crank[D.box_off + i] = __ldcg(P.crank + oc + D.run_lo + i);
                    if (P.class_aware) s_clab[D.box_off + i] = __ldcg(P.clabel + oc + D.run_lo + i);
                }
            }
        }
        __syncthreads();
        // ---- groups of K7_PAIRS_GW lanes take the CTA's boxes one at a time (all control flow is uniform in a group)
        s_mb[tid] = mb;
        s_mri[tid] = ri;
        s_mlab[tid] = mlabel;
        __syncthreads();
        constexpr int GW = K7_PAIRS_GW, NG = 32 / GW, GBUF = K7_PAIRS_EBUF / ((K7_PAIRS_THREADS / 32) * NG);
        const int lane = tid & 31, warp = tid >> 5, gl = lane & (GW - 1), grp = lane / GW;
        const unsigned int gmask = (GW == 32) ? 0xffffffffu : (((1u << GW) - 1u) << (grp * GW));
        const unsigned int lt_mask = ((1u << lane) - 1u) & gmask;     // the group's lanes below mine
        unsigned int* gbuf = s_edge + (warp * NG + grp) * GBUF;
        uint2* grun = s_grun + (warp * NG + grp) * GW;                // the group's runs: {start index, inclusive prefix}
        unsigned int* edges = P.edges + (size_t)b * P.ecap;
        int fill = 0;                                                 // group-uniform
        auto flush = [&]() {
            int base = 0;
            __syncwarp(gmask);            // the lanes' buffered edges are visible
            if (gl == 0) base = atomicAdd(P.nedge + b, fill);
            base = __shfl_sync(gmask, base, grp * GW);
            for (int i = gl; i < fill; i += GW)
                if (base + i < P.ecap) edges[base + i] = gbuf[i];
            __syncwarp(gmask);
            fill = 0;
        };
        for (int it = 0; it < GW; ++it) {   // 32 boxes per warp, NG at a time
            const int q = (warp << 5) + it * NG + grp;
            const int bri = s_mri[q];
            if (bri < 0) continue;          // group-uniform
            const float4 bb = s_mb[q];
            const int blabel = s_mlab[q];
            const K7Box me2 = {bb.x, bb.y, bb.z, bb.w};
            const float bw = K7_FSUB(bb.z, bb.x), bh = K7_FSUB(bb.w, bb.y), barea = K7_FMUL(bw, bh);
            const float bca = K7_FMUL(c_flt, barea);
            const float bcx = 0.5f * (bb.x + bb.z), bcy = 0.5f * (bb.y + bb.w);
            const float brx = kq * bw + 1e-2f, bry = kq * bh + 1e-2f;
            const int bclo = k7_grid_class(fmaxf(barea * tl * (1.0f - 1e-5f), 1e-30f));
            const int bchi = k7_grid_class(fminf(barea / tl * (1.0f + 1e-5f), 3.0e38f));
            int R = 0;                      // row runs of the neighbourhood: one per (class, grid row)
            for (int cl = bclo; cl <= bchi; ++cl) {
                const int gy = s_g.gy[cl];
                const float ich = s_g.inv_ch[cl];
                const int y0 = max(0, min(gy - 1, (int)floorf((bcy - bry - s_g.oy) * ich))), y1 = max(0, min(gy - 1, (int)floorf((bcy + bry - s_g.oy) * ich)));
                R += y1 - y0 + 1;
            }
            int count = 0;
            for (int r0 = 0; r0 < R; r0 += GW) {
                // ---- my row run: cells (iy, x0 .. x1) of one class are consecutive in the cell order
                int lo = BIG, hi = 0, start = 0;
                {
                    int rr = r0 + gl;
                    for (int cl = bclo; cl <= bchi; ++cl) {
                        const int gx = s_g.gx[cl], gy = s_g.gy[cl];
                        const float icw = s_g.inv_cw[cl], ich = s_g.inv_ch[cl];
                        const int y0 = max(0, min(gy - 1, (int)floorf((bcy - bry - s_g.oy) * ich))), y1 = max(0, min(gy - 1, (int)floorf((bcy + bry - s_g.oy) * ich)));
                        const int nrows = y1 - y0 + 1;
                        if (rr >= nrows) {
                            rr -= nrows;
                            continue;
                        }
                        const int iy = y0 + rr;
                        const int x0 = max(0, min(gx - 1, (int)floorf((bcx - brx - s_g.ox) * icw))), x1 = max(0, min(gx - 1, (int)floorf((bcx + brx - s_g.ox) * icw)));
                        const K7PairsCls D = staged_tab ? s_cls[cl - CLO] : K7PairsCls{};
                        for (int ix = x0; ix <= x1; ++ix) {
                            const uint2 se = staged_tab ? s_tab[D.tab_off + (iy - D.row0) * gx + ix] : __ldcg(tab + s_g.off[cl] + iy * gx + ix);
                            if (se.y) {
                                lo = min(lo, P.cap - (int)se.x);
                                hi = max(hi, (int)se.y);
                            }
                        }
                        start = staged ? D.box_off + (lo - D.run_lo) : lo;
                        break;
                    }
                }
                const int len = max(0, hi - lo);
                int inc = len;
#pragma unroll
                for (int o = 1; o < GW; o <<= 1) {
                    const int t = __shfl_up_sync(gmask, inc, o, GW);
                    if (gl >= o) inc += t;
                }
                const int T = __shfl_sync(gmask, inc, GW - 1, GW);
                if (T == 0) continue;       // group-uniform
                __syncwarp(gmask);          // the previous pass is done with the run table
                grun[gl] = make_uint2((unsigned int)start, (unsigned int)inc);
                __syncwarp(gmask);
                // ---- the T candidates of this pass, cut into GW equal parts
                int t = (int)(((long long)T * gl) / GW);
                const int t_end = (int)(((long long)T * (gl + 1)) / GW);
                int r = 0;
                while (r < GW - 1 && (int)grun[r].y <= t) ++r;
                int run_end = (int)grun[r].y;
                int idx = (int)grun[r].x + (t - (r ? (int)grun[r - 1].y : 0));
                const int steps = (T + GW - 1) / GW;
                for (int sidx = 0; sidx < steps; ++sidx) {   // group-uniform trip count
                    bool hit = false;
                    int rj = 0;
                    if (t < t_end) {
                        float4 ob;
                        int lab = 0;
                        if (staged) {
                            rj = s_crank[idx];
                            ob = s_cbox[idx];
                            if (P.class_aware) lab = s_clab[idx];
                        } else {
                            rj = __ldcg(P.crank + oc + idx);
                            ob = __ldcg(P.cbox + oc + idx);
                            if (P.class_aware) lab = __ldcg(P.clabel + oc + idx);
                        }
                        // only boxes visited before mine can suppress it; class-aware: other classes never do
                        hit = rj < bri && lab == blabel &&
                              k7_nms_decide(me2, barea, bca, ob, K7_FMUL(c_flt, k7_area(ob.x, ob.y, ob.z, ob.w)), thr);
                        ++t;
                        ++idx;
                        if (t == run_end && t < t_end) {   // on to the next run that has boxes
                            do ++r; while (r < GW - 1 && (int)grun[r].y <= t);
                            run_end = (int)grun[r].y;
                            idx = (int)grun[r].x;
                        }
                    }
                    const unsigned int m = __ballot_sync(gmask, hit);
                    if (m == 0u) continue;
                    const int c = __popc(m);
                    if (fill + c > GBUF) flush();
                    if (hit) gbuf[fill + __popc(m & lt_mask)] = ((unsigned int)bri << 16) | (unsigned int)rj;
                    fill += c;
                    count += c;
                }
            }
            if (gl == 0) P.pend[oc + bri] = count;
        }
        __syncwarp(gmask);
        if (fill > 0) flush();
    }
}

// ---- pair tests for images with few boxes (n <= brute_n <= 2048, the usual detector image: a few hundred
// survivors).  All n(n-1)/2 pairs are cheaper than building cells: 345 k tests for 831 boxes, spread over 13 CTAs.
// A CTA takes 64 consecutive ranks; the boxes ranked before its last row wait in shared memory.  A thread owns a
// row and every fourth 64-candidate WORD of its suppression bit row (a warp's lanes share the candidate: one
// broadcast load per test); the word is built in a register, so a test is ~16 instructions and no atomic - the
// first version appended every hit to a buffer (one shared-memory atomic per warp and test with a hit, and a
// global one once 2048 edges were in: 67 us for 64 x 831 boxes, ten times the arithmetic).  The bit rows then
// become edges of the format k7_nms_pairs_kernel writes: row counts -> scan -> ONE reservation of the exact size.
// Same filter and exact test as the grid kernel; blocks with many candidates are handed out first.
#define K7_BRUTE_N 1024          // default brute_n
#define K7_BRUTE_THREADS 256
#define K7_BRUTE_ROWS 64
#define K7_BRUTE_PARTS (K7_BRUTE_THREADS / K7_BRUTE_ROWS)
static size_t k7_brute_smem(int brute_n) {
    const size_t ncap = (size_t)(brute_n + K7_BRUTE_ROWS - 1) / K7_BRUTE_ROWS * K7_BRUTE_ROWS;
    return ncap * (sizeof(float4) + sizeof(float) + sizeof(int)) + (size_t)K7_BRUTE_ROWS * (ncap / 64 + 1) * sizeof(unsigned long long);
}
__global__ void __launch_bounds__(K7_BRUTE_THREADS)
k7_nms_brute_kernel(const __grid_constant__ K7NmsArgs P, int B) {
    extern __shared__ __align__(16) unsigned char dyn_raw[];   // k7_brute_smem(brute_n) bytes
    const int ncap = (P.brute_n + K7_BRUTE_ROWS - 1) / K7_BRUTE_ROWS * K7_BRUTE_ROWS;
    const int pitch = ncap / 64 + 1;                           // u64 words per bit row (+1: rows fall on different banks)
    float4* s_box = reinterpret_cast<float4*>(dyn_raw);
    unsigned long long* s_mask = reinterpret_cast<unsigned long long*>(s_box + ncap);
    float* s_ca = reinterpret_cast<float*>(s_mask + K7_BRUTE_ROWS * pitch);
    int* s_lab = reinterpret_cast<int*>(s_ca + ncap);
    __shared__ int s_cnt[K7_BRUTE_ROWS], s_off[K7_BRUTE_ROWS];
    __shared__ int s_ebase;
    const int tid = threadIdx.x, lane = tid & 31;
    k7_pdl_trigger();
    k7_pdl_wait();
    const int per_img = ncap / K7_BRUTE_ROWS;
    const float thr = P.thr;
    const float c_flt = thr / (1.0f + thr) * (1.0f - 1e-5f);
    for (int item = blockIdx.x; item < per_img * B; item += gridDim.x) {   // block-uniform
        const int blk = per_img - 1 - item / B, b = item % B;              // the longest rows first
        const int n = P.nsel[b];
        if (n < 0 || P.surv_count[b] > P.brute_n) continue;
        const int i0 = blk * K7_BRUTE_ROWS;
        if (i0 >= n) continue;
        const size_t ob = (size_t)b * P.A, oc = (size_t)b * P.cap;
        const int nj = min(n, i0 + K7_BRUTE_ROWS);
        __syncthreads();   // the previous item is done with shared memory
        for (int j = tid; j < ((nj + 63) & ~63); j += K7_BRUTE_THREADS) {
            const float4 bx = j < nj ? __ldcg(P.rbox + oc + j) : make_float4(0.f, 0.f, 0.f, 0.f);   // padding: never hits
            s_box[j] = bx;
            s_ca[j] = K7_FMUL(c_flt, k7_area(bx.x, bx.y, bx.z, bx.w));
            if (P.class_aware) s_lab[j] = j < nj ? __ldg(P.label_ws + ob + (uint32_t)(__ldcg(P.keys1 + ob + j) & 0xffffffffull)) : -1;
        }
        __syncthreads();
        const int r = tid & (K7_BRUTE_ROWS - 1), ri = i0 + r, part = tid / K7_BRUTE_ROWS;
        const bool row = ri < n;
        const int nw = blk + 1;   // words 0 .. blk hold the candidates ranked before the CTA's rows
        {
            const float4 mb = row ? s_box[ri] : make_float4(0.f, 0.f, 0.f, 0.f);
            const K7Box me = {mb.x, mb.y, mb.z, mb.w};
            const float marea = k7_area(mb.x, mb.y, mb.z, mb.w), mca = row ? s_ca[ri] : 0.0f;
            const int mlabel = (row && P.class_aware) ? s_lab[ri] : 0;
            // Fast path per candidate: no branch, no division.  `yes` = suppressed for sure (the sufficient condition of
            // k7_nms_decide), `maybe` = passed the necessary filter only: those few (within 2.2e-5 of the threshold, or
            // overflowed) get the exact test afterwards.  32 candidates per register word with compile-time bit positions.
            for (int w = part; w < nw; w += K7_BRUTE_PARTS) {
                unsigned int yes[2] = {0u, 0u}, maybe[2] = {0u, 0u};
                const int j0 = w << 6;
#pragma unroll
                for (int hw = 0; hw < 2; ++hw) {
                    const int jb = j0 + 32 * hw;
                    if (jb >= nj) break;   // block-uniform
#pragma unroll 8
                    for (int k = 0; k < 32; ++k) {
                        const float4 cb = s_box[jb + k];                       // one address per warp: a broadcast
                        const float cca = s_ca[jb + k];                        // (beyond nj: zero boxes, never hit: w <= 0)
                        const float iw = K7_FSUB(fminf(me.x2, cb.z), fmaxf(me.x1, cb.x));
                        const float ih = K7_FSUB(fminf(me.y2, cb.w), fmaxf(me.y1, cb.y));
                        const float inter = K7_FMUL(iw, ih), sm = K7_FADD(mca, cca);
                        const bool cand = (iw > 0.0f) & (ih > 0.0f) & (inter > sm);
                        const bool sure = cand & (inter > K7_FMUL(sm, 1.0000221f)) & (inter <= 3.0e38f);
                        yes[hw] |= sure ? (1u << k) : 0u;
                        maybe[hw] |= (cand & !sure) ? (1u << k) : 0u;
                    }
                }
                unsigned long long word = ((unsigned long long)yes[1] << 32) | yes[0];
                unsigned long long mb2 = ((unsigned long long)maybe[1] << 32) | maybe[0];
                while (mb2) {   // rare
                    const int k = __ffsll((long long)mb2) - 1;
                    mb2 &= mb2 - 1ull;
                    const float4 cb = s_box[j0 + k];
                    if (k7_nms_suppress_exact(me, marea, cb, k7_area(cb.x, cb.y, cb.z, cb.w), thr)) word |= 1ull << k;
                }
                if (w == blk) {   // the diagonal word: only boxes ranked before mine
                    word &= (r == 0) ? 0ull : (~0ull >> (64 - r));
                }
                if (P.class_aware) {
                    unsigned long long t = word;
                    while (t) {
                        const int k = __ffsll((long long)t) - 1;
                        t &= t - 1ull;
                        if (s_lab[j0 + k] != mlabel) word &= ~(1ull << k);
                    }
                }
                s_mask[r * pitch + w] = row ? word : 0ull;
            }
        }
        __syncthreads();
        // ---- bit rows -> edges: counts per row, scan, one reservation
        if (tid < K7_BRUTE_ROWS) {
            int c = 0;
            for (int w = 0; w < nw; ++w) c += __popcll(s_mask[tid * pitch + w]);
            s_cnt[tid] = c;
            if (i0 + tid < n) P.pend[oc + i0 + tid] = c;
            int inc = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            s_off[tid] = inc - c;   // exclusive within the warp; the second warp adds the first one's total below
        }
        __syncthreads();
        const int half = s_off[31] + s_cnt[31];
        const int total = half + s_off[63] + s_cnt[63];
        if (total == 0) continue;   // block-uniform
        if (tid == 0) s_ebase = atomicAdd(P.nedge + b, total);
        __syncthreads();
        unsigned int* edges = P.edges + (size_t)b * P.ecap;
        {
            int at = s_ebase + s_off[r] + (r >= 32 ? half : 0);
            for (int w = 0; w < nw; ++w) {
                unsigned long long word = s_mask[r * pitch + w];
                if ((w & (K7_BRUTE_PARTS - 1)) != part) {   // another thread's word
                    at += __popcll(word);
                    continue;
                }
                while (word) {
                    const int k = __ffsll((long long)word) - 1;
                    word &= word - 1ull;
                    if (at < P.ecap) edges[at] = ((unsigned int)ri << 16) | (unsigned int)((w << 6) | k);
                    ++at;
                }
            }
        }
    }
}

// ---- Greedy NMS as a fixed point over the edges, one CTA per image.  Box i is REMOVED as soon as one of
// its suppressors (edges (i, j): j ranks before i and overlaps it above thr) is known KEPT, and KEPT as
// soon as all of them are known removed; the undecided box of lowest rank can always be decided (all
// its suppressors rank before it, hence are decided), so the iteration ends, and by induction over the
// rank its result is exactly the sequential greedy answer of torchvision.ops.nms.  The kept / removed
// sets only grow, so threads may read them while others update them; a round costs two barriers.
// Rounds = longest dependency chain (2-3 on the bench inputs).  An edge whose suppressor was removed is
// dead; the box's count of live edges drops, and at zero the box is kept.
#define K7_RESOLVE_THREADS 1024
#define K7_RESOLVE_EDGES 32768      // edges staged in shared memory; longer lists are walked in global memory
#define K7_RESOLVE_SMEM_NEED (K7_RESOLVE_EDGES * sizeof(unsigned int) + K7_SORT_MAX * sizeof(int))
#define K7_RESOLVE_SMEM (K7_RESOLVE_SMEM_NEED > sizeof(K7SortNmsSmem) ? K7_RESOLVE_SMEM_NEED : sizeof(K7SortNmsSmem))
__device__ __forceinline__ void k7_set_bit32(unsigned int* set, int p) { atomicOr(set + (p >> 5), 1u << (p & 31)); }

__global__ void __launch_bounds__(K7_RESOLVE_THREADS)
k7_nms_resolve_kernel(const __grid_constant__ K7NmsArgs P, int max_det, uint64_t* __restrict__ kept_key,
                      float4* __restrict__ kept_box, int32_t* __restrict__ kept_count, int32_t* __restrict__ counts_out,
                      int64_t* __restrict__ keep_idx_out) {
    constexpr int WMAX = K7_SORT_MAX / 32;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    unsigned int* s_e = reinterpret_cast<unsigned int*>(smem_raw);
    int* s_pend = reinterpret_cast<int*>(s_e + K7_RESOLVE_EDGES);
    __shared__ unsigned int s_kept[WMAX], s_rem[WMAX];
    __shared__ int s_prefix[WMAX + 1];
    __shared__ int s_open;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    k7_pdl_trigger();
    k7_pdl_wait();
    const int n = P.nsel[b];
    const int E = n >= 0 ? P.nedge[b] : 0;
    const size_t ob = (size_t)b * P.A, oc = (size_t)b * P.cap;
    if (n < 0 || E > P.ecap) {   // too many survivors / edges for the multi-SM path, or a tiny threshold
        k7_fallback_image(b, P.A, P.thr, P.pre_topk, max_det, P.surv_count, kept_count, P.keys0, P.keys1, P.box_ws, kept_key,
                          kept_box, counts_out, P.class_aware ? P.label_ws : nullptr, *reinterpret_cast<K7SortNmsSmem*>(smem_raw));
        if (keep_idx_out) {
            __syncthreads();
            const int kc = kept_count[b];
            for (int k = tid; k < kc; k += K7_RESOLVE_THREADS) keep_idx_out[k] = (int64_t)(uint32_t)(kept_key[ob + k] & 0xffffffffull);
        }
        return;
    }
    const int nw = (n + 31) >> 5;
    for (int w = tid; w < WMAX; w += K7_RESOLVE_THREADS) s_kept[w] = s_rem[w] = 0u;
    if (tid == 0) s_open = 0;
    __syncthreads();
    // boxes without an edge are kept at once; the others wait with their count of live edges
    int open_mine = 0;
    for (int r0 = 0; r0 < n; r0 += K7_RESOLVE_THREADS) {
        const int r = r0 + tid;
        const int pd = (r < n) ? __ldcg(P.pend + oc + r) : 0;
        if (r < n) s_pend[r] = pd;
        const unsigned int k0 = __ballot_sync(0xffffffffu, r < n && pd == 0);
        if (lane == 0 && (r0 + (warp << 5)) < n) s_kept[(r0 >> 5) + warp] = k0;
        open_mine += (r < n && pd > 0) ? 1 : 0;
    }
    if (open_mine) atomicAdd(&s_open, open_mine);
    const bool staged = E <= K7_RESOLVE_EDGES;
    const unsigned int* ge = P.edges + (size_t)b * P.ecap;
    if (staged)
        for (int i = tid; i < E; i += K7_RESOLVE_THREADS) s_e[i] = __ldcg(ge + i);
    for (int round = 0; round <= n + 1; ++round) {   // every round decides at least the lowest-ranked open box
        __syncthreads();
        const int open = s_open;   // read by everybody between two barriers
        __syncthreads();
        if (open <= 0) break;
        for (int i = tid; i < E; i += K7_RESOLVE_THREADS) {
            const unsigned int e = staged ? s_e[i] : __ldcg(ge + i);
            if (e == 0xffffffffu) continue;                          // dead (only staged lists are marked)
            const int ri = (int)(e >> 16), rj = (int)(e & 0xffffu);
            const unsigned int bit_i = 1u << (ri & 31);
            if ((*((volatile unsigned int*)&s_kept[ri >> 5]) | *((volatile unsigned int*)&s_rem[ri >> 5])) & bit_i) continue;   // decided
            const unsigned int bit_j = 1u << (rj & 31);
            if (*((volatile unsigned int*)&s_kept[rj >> 5]) & bit_j) {
                if (!(atomicOr(&s_rem[ri >> 5], bit_i) & bit_i)) atomicSub(&s_open, 1);     // first to remove it
            } else if (staged && (*((volatile unsigned int*)&s_rem[rj >> 5]) & bit_j)) {
                s_e[i] = 0xffffffffu;                                // this suppressor is gone
                if (atomicSub(&s_pend[ri], 1) == 1) {                // its last live edge: nothing that is kept suppresses it
                    atomicOr(&s_kept[ri >> 5], bit_i);
                    atomicSub(&s_open, 1);
                }
            }
        }
        if (!staged) {
            // A long edge list stays in global memory and cannot be marked: recount instead.  A box is kept when
            // every one of its suppressors is removed.
            __syncthreads();
            for (int r = tid; r < n; r += K7_RESOLVE_THREADS) s_pend[r] = 0;
            __syncthreads();
            for (int i = tid; i < E; i += K7_RESOLVE_THREADS) {
                const unsigned int e = __ldcg(ge + i);
                const int ri = (int)(e >> 16), rj = (int)(e & 0xffffu);
                if (!((s_rem[rj >> 5] >> (rj & 31)) & 1u)) atomicAdd(&s_pend[ri], 1);   // a suppressor that is not removed (yet)
            }
            __syncthreads();
            for (int r = tid; r < n; r += K7_RESOLVE_THREADS) {
                const unsigned int bit = 1u << (r & 31);
                if (!((s_kept[r >> 5] | s_rem[r >> 5]) & bit) && s_pend[r] == 0) {
                    atomicOr(&s_kept[r >> 5], bit);
                    atomicSub(&s_open, 1);
                }
            }
        }
    }
    // kept set is in rank (= visiting) order: compaction
    if (tid < 32) {   // exclusive prefix of the popcounts of s_kept[0..nw)
        int carry = 0;
        for (int w0 = 0; w0 < nw; w0 += 32) {
            const int w = w0 + tid;
            const int v = (w < nw) ? __popc(s_kept[w]) : 0;
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
    for (int r = tid; r < n; r += K7_RESOLVE_THREADS) {
        if (!((s_kept[r >> 5] >> (r & 31)) & 1u)) continue;
        const int pos = s_prefix[r >> 5] + __popc(s_kept[r >> 5] & ((1u << (r & 31)) - 1u));
        if (pos >= budget) continue;
        const uint64_t key = __ldcg(P.keys1 + ob + r);
        kept_key[ob + pos] = key;
        kept_box[ob + pos] = __ldcg(P.rbox + oc + r);
        if (keep_idx_out) keep_idx_out[pos] = (int64_t)(uint32_t)(key & 0xffffffffull);
    }
    if (tid == 0) {
        const int kc = min(s_prefix[nw], budget);
        kept_count[b] = kc;
        if (counts_out) {
            counts_out[2 * b + 0] = P.surv_count[b];
            counts_out[2 * b + 1] = kc;
        }
    }
}

// exclusive popcount prefix of the survivor bitmap (only when the caller wants survivor ranks)
__global__ void __launch_bounds__(K7_NMS_THREADS)
k7_rank_prefix_kernel(int W32, const uint32_t* __restrict__ bits, int32_t* __restrict__ wprefix) {
    __shared__ int s_w[32];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    k7_pdl_wait();
    const int chunk = (W32 + K7_NMS_THREADS - 1) / K7_NMS_THREADS;
    const int w_lo = min(tid * chunk, W32), w_hi = min(w_lo + chunk, W32);
    const uint32_t* brow = bits + (size_t)b * W32;
    int cnt = 0;
    for (int w = w_lo; w < w_hi; ++w) cnt += __popc(brow[w]);
    int inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    int run = inc - cnt;
    for (int k = 0; k < warp; ++k) run += s_w[k];
    for (int w = w_lo; w < w_hi; ++w) {
        wprefix[(size_t)b * W32 + w] = run;
        run += __popc(brow[w]);
    }
}

// ------------------------------------------------------------------ host: the NMS chain
static K7NmsArgs k7_nms_args(const K7DetectWs& w, int A, float thr, int pre_topk, bool class_aware) {
    K7NmsArgs P;
    memset(&P, 0, sizeof(P));
    P.A = A; P.cap = w.cap; P.pre_topk = pre_topk; P.class_aware = class_aware ? 1 : 0; P.thr = thr;
    P.surv_count = w.surv_count; P.keys0 = w.keys0; P.keysC = w.keysC; P.keys1 = w.keys1; P.box_ws = w.box;
    P.label_ws = w.label; P.rbox = w.rbox; P.bounds = w.bounds; P.nsel = w.nsel; P.ccnt = w.ccnt; P.ckeyC = w.ckeyC;
    P.cbox = w.cbox; P.crank = w.crank; P.clabel = w.clabel; P.ctab = w.ctab; P.pend = w.pend; P.edges = w.edges;
    P.nedge = w.nedge; P.ecap = w.ecap;
    { static int sn = -2; if (sn == -2) { const char* e = getenv("K7_NMS_BRUTE_N"); sn = e ? atoi(e) : K7_BRUTE_N; if (sn > K7_NMS_CH) sn = K7_NMS_CH; } P.brute_n = sn; }
    static_assert(sizeof(K7Grid) <= 320 && sizeof(K7Grid) % 4 == 0 && sizeof(K7Grid) / 4 <= K7_PAIRS_THREADS, "K7Grid is copied by one thread per word");
    P.ginfo = static_cast<K7Grid*>(w.ginfo);
    return P;
}

static int k7_nms_chain(const K7NmsArgs& P, int B, int max_det, uint64_t* kept_key, float4* kept_box, int32_t* kept_count,
                        int32_t* counts, int64_t* keep_idx_out, cudaStream_t st) {
    const int G = P.cap / K7_NMS_CH;
    const float tl = P.thr * (1.0f - 1e-4f);
    const float kq = P.thr >= K7_GRID_MIN_THR ? fmaxf(1.0f - tl, 0.5f / tl - 0.5f) * (1.0f + 1e-4f) : 0.0f;
    const int sms = k7_sm_count();
    auto grid1 = [&](long long items, int per_sm) { return dim3((unsigned)(items < (long long)per_sm * sms ? (items > 0 ? items : 1) : (long long)per_sm * sms)); };
    k7_pdl_site() = 0;
    K7_CUDA(k7_launch_pdl(k7_nms_chunksort_kernel, grid1((long long)G * B, 2), dim3(K7_NMS_THREADS), 0, st, P, B));   // also sets nsel
    k7_pdl_site() = 1;
    K7_CUDA(k7_launch_pdl(k7_nms_rank_kernel<false>, grid1(B, 2), dim3(K7_NMS_THREADS), 0, st, P, B));
    if (G > 1) {
        const size_t smem_rank = (size_t)P.cap * sizeof(unsigned long long);   // all keys of a multi-chunk image
        K7_CUDA(cudaFuncSetAttribute(k7_nms_rank_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_rank));
        K7_CUDA(k7_launch_pdl(k7_nms_rank_kernel<true>, grid1((long long)G * B, 1), dim3(K7_NMS_THREADS), smem_rank, st, P, B));
        K7_CUDA(k7_launch_pdl(k7_nms_cellorder_kernel, grid1((long long)G * B, 2), dim3(K7_NMS_THREADS), 0, st, P, B));
    }
    K7_CUDA(cudaFuncSetAttribute(k7_nms_pairs_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K7_PAIRS_SMEM));
    K7_CUDA(k7_launch_pdl(k7_nms_pairs_kernel, grid1((long long)(P.cap / K7_PAIRS_THREADS) * B, 3), dim3(K7_PAIRS_THREADS), K7_PAIRS_SMEM, st, P, kq, B));
    if (P.brute_n > 0) {
        const size_t smem_b = k7_brute_smem(P.brute_n);
        K7_CUDA(cudaFuncSetAttribute(k7_nms_brute_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_b));
        K7_CUDA(k7_launch_pdl(k7_nms_brute_kernel, grid1((long long)((P.brute_n + K7_BRUTE_ROWS - 1) / K7_BRUTE_ROWS) * B, 6),
                              dim3(K7_BRUTE_THREADS), smem_b, st, P, B));
    }
    K7_CUDA(cudaFuncSetAttribute(k7_nms_resolve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K7_RESOLVE_SMEM));
    k7_pdl_site() = 2;
    K7_CUDA(k7_launch_pdl(k7_nms_resolve_kernel, dim3(B), dim3(K7_RESOLVE_THREADS), K7_RESOLVE_SMEM, st, P, max_det, kept_key,
                          kept_box, kept_count, counts, keep_idx_out));
    return K7_OK;
}

// ------------------------------------------------------------------ gather
__global__ void __launch_bounds__(256)
k7_gather_kernel(int A, int W32, int cap, const int32_t* __restrict__ kept_count, const uint64_t* __restrict__ kept_key,
                 const float4* __restrict__ kept_box, const int32_t* __restrict__ label_ws, const uint32_t* __restrict__ bits,
                 const int32_t* __restrict__ wprefix, float* __restrict__ scores, int64_t* __restrict__ labels,
                 float4* __restrict__ boxes, int32_t* __restrict__ anchor_ids, int32_t* __restrict__ ranks) {
    const int b = blockIdx.y;
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    k7_pdl_wait();      // k7_resolve_kernel's kept lists
    const int kc = min(kept_count[b], cap);
    if (k >= kc) return;
    const size_t ob = (size_t)b * A;
    const uint64_t key = kept_key[ob + k];
    const int a = (int)(uint32_t)(key & 0xffffffffull);
    const size_t o = (size_t)b * cap + k;
    scores[o] = k7_key_score(key);
    labels[o] = (int64_t)label_ws[ob + a];
    boxes[o] = kept_box[ob + k];
    if (anchor_ids) anchor_ids[o] = a;
    if (ranks) {
        const size_t wi = (size_t)b * W32 + (a >> 5);
        ranks[o] = wprefix[wi] + __popc(bits[wi] & ((1u << (a & 31)) - 1u));
    }
}

// ------------------------------------------------------------------ host: k7_detect
static int k7_detect_run(const K7AnchorSpec* spec, const float* anchors, int box_order, const float* cls,
                         const float* const* levels, const float* reg, int B, int A, int C, const K7DetectParams* params,
                         int32_t* counts, void* workspace, size_t workspace_bytes, void* stream) {
    K7_REQUIRE((cls || levels) && reg && params && workspace, "k7_detect: null argument");
    K7_REQUIRE(spec || anchors, "k7_detect: need an anchor spec or an anchor tensor");
    K7_REQUIRE(B > 0 && A > 0 && C > 0, "k7_detect: bad shape B=%d A=%d C=%d", B, A, C);
    K7_REQUIRE(B <= 65535, "k7_detect: B > 65535 not supported");
    K7_REQUIRE(params->iou_thr >= 0.0f, "k7_detect: iou_thr must be >= 0");
    K7_REQUIRE(params->decode.variant == K7_DECODE_RETINA || params->decode.variant == K7_DECODE_EFFDET,
               "k7_detect: bad decode variant");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(cls) & 15) == 0 && (reinterpret_cast<uintptr_t>(reg) & 15) == 0,
               "k7_detect: cls and reg must be 16-byte aligned");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "k7_detect: workspace must be 256-byte aligned");
    if (spec) {
        int rc = k7_check_spec(spec, A);
        if (rc) return rc;
        box_order = spec->box_order;
    } else {
        K7_REQUIRE((reinterpret_cast<uintptr_t>(anchors) & 15) == 0, "k7_detect: anchors must be 16-byte aligned");
        K7_REQUIRE(box_order == K7_BOX_XYXY || box_order == K7_BOX_YXYX, "k7_detect: bad box_order");
    }
    const K7DetectWs w = k7_detect_ws(workspace, B, A);
    if (workspace_bytes < w.bytes) {
        k7_set_error("k7_detect: workspace %zu < %zu bytes", workspace_bytes, w.bytes);
        return K7_EWORKSPACE;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // counters, bounds, the survivor bitmap and the cell table are contiguous at the start of the workspace
    K7_CUDA(cudaMemsetAsync(w.surv_count, 0, w.zero_bytes, st));
    static const K7AnchorSpec kEmpty = {};
    const K7AnchorSpec& sp = spec ? *spec : kEmpty;
    float report_thr = params->score_thr;
    if (params->from_logits) {
        // logit(thr) lowered by a margin: the exact float32 sigmoid decides in the kernel
        const double t = (double)params->score_thr;
        if (t <= 0.0) report_thr = -INFINITY;
        else if (t >= 1.0) report_thr = INFINITY;   // sigmoid <= 1: nothing can pass a strict '>'
        else {
            const double z = log(t / (1.0 - t));
            report_thr = (float)(z - 1e-3 * (1.0 + fabs(z)));
        }
    }
    if (levels) {
        K7ScoreLevels LV;
        memset(&LV, 0, sizeof(LV));
        int tiles = 0, sblocks = 0;
        for (int l = 0; l < spec->num_levels; ++l) {
            const int S = spec->fh[l] * spec->fw[l];
            K7_REQUIRE(S == 0 || (levels[l] && (reinterpret_cast<uintptr_t>(levels[l]) & 15) == 0),
                       "k7_detect_levels: level %d: null or unaligned tensor", l);
            K7_REQUIRE(S < (1 << 27) && spec->cell_anchors <= 16, "k7_detect_levels: level %d: map too large", l);
            LV.cls[l] = levels[l];
            LV.S[l] = S;
            LV.blk_off[l] = tiles;
            LV.sblk_off[l] = sblocks;
            if ((S & 3) == 0) tiles += (spec->cell_anchors * (S / 4) + 255) / 256;
            else sblocks += (spec->cell_anchors * S + 255) / 256;
        }
        for (int l = spec->num_levels; l <= K7_MAX_LEVELS; ++l) {
            LV.blk_off[l] = tiles;
            LV.sblk_off[l] = sblocks;
        }
        K7_REQUIRE((long long)tiles * B < (1ll << 31) && B < (1 << 24), "k7_detect_levels: batch too large");
        if (tiles > 0) {
            const int num_tiles = tiles * B;
            const int grid = num_tiles < 2 * k7_sm_count() ? num_tiles : 2 * k7_sm_count();
            auto kern = params->from_logits ? k7_score_levels_kernel<true> : k7_score_levels_kernel<false>;
            K7_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K7_LV_SMEM));
            kern<<<grid, 256, K7_LV_SMEM, st>>>(sp, LV, num_tiles, reinterpret_cast<const float4*>(reg), A, C, report_thr, *params,
                                                w.bits, w.W32, w.box, w.label, w.surv_count, w.keys0);
            K7_LAUNCH_CHECK();
        }
        if (sblocks > 0) {
            dim3 grid(sblocks, B);
            auto kern = params->from_logits ? k7_score_levels_scalar_kernel<true> : k7_score_levels_scalar_kernel<false>;
            kern<<<grid, 256, 0, st>>>(sp, LV, reinterpret_cast<const float4*>(reg), A, C, report_thr, *params, w.bits, w.W32,
                                       w.box, w.label, w.surv_count, w.keys0);
            K7_LAUNCH_CHECK();
        }
    } else {
        K7ScoreGeom GS;
        int rc = k7_score_geometry(B, A, C, &GS);
        if (rc) return rc;
#define K7_SPIPE_LAUNCH(SP, LG)                                                                                          \
    do {                                                                                                                 \
        auto kern = k7_score_pipe_kernel<SP, LG>;                                                                        \
        K7_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GS.smem_bytes));            \
        kern<<<GS.grid, (GS.cw + 1) * 32, GS.smem_bytes, st>>>(sp, anchors, box_order, cls,                              \
                                                               reinterpret_cast<const float4*>(reg), A, C, report_thr,  \
                                                               *params, GS, w.bits, w.W32, w.box, w.label, w.surv_count, w.keys0);    \
    } while (0)
        if (params->from_logits) { if (spec) K7_SPIPE_LAUNCH(true, true); else K7_SPIPE_LAUNCH(false, true); }
        else                     { if (spec) K7_SPIPE_LAUNCH(true, false); else K7_SPIPE_LAUNCH(false, false); }
#undef K7_SPIPE_LAUNCH
        K7_LAUNCH_CHECK();
    }

    // NMS (every kernel decides per image on the device; PDL links: each starts when its predecessor drains)
    int rc2;
    K7_REQUIRE(params->nms_chain >= 0 && params->nms_chain <= 2, "k7_detect: bad nms_chain %d", params->nms_chain);
    const bool per_image = params->nms_chain == K7_NMS_CHAIN_AUTO ? w.use_v1 : params->nms_chain == K7_NMS_CHAIN_PER_IMAGE;
    if (per_image)
        rc2 = k7v1::chain(B, A, params->iou_thr, params->pre_nms_topk, params->max_detections, params->class_aware != 0,
                          w.surv_count, w.keys0, w.keys1, w.box, w.label, w.nsel, w.kept_key, w.kept_box, w.kept_count, counts,
                          w.v1, st);
    else
        rc2 = k7_nms_chain(k7_nms_args(w, A, params->iou_thr, params->pre_nms_topk, params->class_aware != 0), B,
                           params->max_detections, w.kept_key, w.kept_box, w.kept_count, counts, nullptr, st);
    if (rc2) return rc2;
    K7_CUDA(cudaGetLastError());
    return K7_OK;
}

extern "C" int k7_detect(const K7AnchorSpec* spec, const float* anchors, int box_order, const float* cls, const float* reg,
                         int B, int A, int C, const K7DetectParams* params, int32_t* counts, void* workspace,
                         size_t workspace_bytes, void* stream) {
    K7_REQUIRE(cls, "k7_detect: null argument");
    return k7_detect_run(spec, anchors, box_order, cls, nullptr, reg, B, A, C, params, counts, workspace, workspace_bytes,
                         stream);
}

extern "C" int k7_detect_levels(const K7AnchorSpec* spec, const float* const* levels, const float* reg, int B, int C,
                                const K7DetectParams* params, int32_t* counts, void* workspace, size_t workspace_bytes,
                                void* stream) {
    K7_REQUIRE(spec && levels, "k7_detect_levels: an anchor spec and the level tensors are required");
    return k7_detect_run(spec, nullptr, spec->box_order, nullptr, levels, reg, B, spec->total, C, params, counts, workspace,
                         workspace_bytes, stream);
}

extern "C" int k7_detect_gather(const void* workspace, size_t workspace_bytes, int B, int A, int C, int cap, float* scores,
                                int64_t* labels, float* boxes, int32_t* anchor_ids, int32_t* survivor_rank, void* stream) {
    (void)C;
    K7_REQUIRE(workspace && scores && labels && boxes, "k7_detect_gather: null argument");
    K7_REQUIRE(B > 0 && A > 0 && cap > 0, "k7_detect_gather: bad shape");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15) == 0, "k7_detect_gather: boxes must be 16-byte aligned");
    const K7DetectWs w = k7_detect_ws(const_cast<void*>(workspace), B, A);
    if (workspace_bytes < w.bytes) {
        k7_set_error("k7_detect_gather: workspace %zu < %zu bytes", workspace_bytes, w.bytes);
        return K7_EWORKSPACE;
    }
    dim3 grid((cap + 255) / 256, B);
    k7_pdl_site() = 3;
    if (survivor_rank)   // the exclusive popcount prefix of the survivor bitmap is only needed for this output
        K7_CUDA(k7_launch_pdl(k7_rank_prefix_kernel, dim3(B), dim3(K7_NMS_THREADS), 0, static_cast<cudaStream_t>(stream), w.W32,
                              w.bits, w.wprefix));
    K7_CUDA(k7_launch_pdl(k7_gather_kernel, grid, dim3(256), 0, static_cast<cudaStream_t>(stream), A, w.W32, cap, w.kept_count,
                          w.kept_key, w.kept_box, w.label, w.bits, w.wprefix, scores, labels, reinterpret_cast<float4*>(boxes),
                          anchor_ids, survivor_rank));
    K7_CUDA(cudaGetLastError());
    return K7_OK;
}

// ------------------------------------------------------------------ stand-alone NMS (torchvision.ops.nms)
// Same kernels as k7_detect with B = 1 and "anchor id" = box index: the workspace is k7_detect's for (1, n).
extern "C" size_t k7_nms_workspace_bytes(int n) { return k7_detect_ws(nullptr, 1, n > 0 ? n : 1, false).bytes; }

__global__ void k7_nms_keys_kernel(const float* __restrict__ scores, int n, uint64_t* __restrict__ keys,
                                   int32_t* __restrict__ n_dev, const int64_t* __restrict__ idxs,
                                   int32_t* __restrict__ labels) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) *n_dev = n;
    if (i < n) {
        keys[i] = ((uint64_t)k7_score_key_desc(scores[i]) << 32) | (uint32_t)i;
        if (idxs) labels[i] = (int32_t)idxs[i];   // only equality matters; category ids fit 32 bits
    }
}

static int k7_nms_run(const float* boxes, const float* scores, const int64_t* idxs, int n, float iou_thr, int64_t* keep,
                      int32_t* count, void* workspace, size_t workspace_bytes, void* stream) {
    K7_REQUIRE(count && workspace, "k7_nms: null argument");
    K7_REQUIRE(n >= 0 && iou_thr >= 0.0f, "k7_nms: bad argument n=%d iou_thr=%f", n, (double)iou_thr);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n == 0) {
        K7_CUDA(cudaMemsetAsync(count, 0, sizeof(int32_t), st));
        return K7_OK;
    }
    K7_REQUIRE(boxes && scores && keep, "k7_nms: null argument");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15) == 0 && (reinterpret_cast<uintptr_t>(workspace) & 255) == 0,
               "k7_nms: boxes must be 16-byte and workspace 256-byte aligned");
    const K7DetectWs w = k7_detect_ws(workspace, 1, n, false);
    if (workspace_bytes < w.bytes) {
        k7_set_error("k7_nms: workspace %zu < %zu bytes", workspace_bytes, w.bytes);
        return K7_EWORKSPACE;
    }
    K7_CUDA(cudaMemsetAsync(w.surv_count, 0, w.zero_bytes, st));
    k7_nms_keys_kernel<<<(n + 255) / 256, 256, 0, st>>>(scores, n, w.keys0, w.surv_count, idxs, w.label);
    K7_LAUNCH_CHECK();
    K7NmsArgs P = k7_nms_args(w, n, iou_thr, 0, idxs != nullptr);
    P.box_ws = reinterpret_cast<const float4*>(boxes);
    int rc = k7_nms_chain(P, 1, 0, w.kept_key, w.kept_box, count, nullptr, keep, st);
    if (rc) return rc;
    K7_CUDA(cudaGetLastError());
    return K7_OK;
}

extern "C" int k7_nms(const float* boxes, const float* scores, int n, float iou_thr, int64_t* keep, int32_t* count,
                      void* workspace, size_t workspace_bytes, void* stream) {
    return k7_nms_run(boxes, scores, nullptr, n, iou_thr, keep, count, workspace, workspace_bytes, stream);
}

extern "C" int k7_batched_nms(const float* boxes, const float* scores, const int64_t* idxs, int n, float iou_thr, int64_t* keep,
                              int32_t* count, void* workspace, size_t workspace_bytes, void* stream) {
    K7_REQUIRE(n == 0 || idxs, "k7_batched_nms: null argument");
    return k7_nms_run(boxes, scores, idxs, n, iou_thr, keep, count, workspace, workspace_bytes, stream);
}

// ------------------------------------------------------------------ k7_nms without a caller workspace
// For callers that may not allocate (graph capture, serving loops): k7_nms with its scratch in a file-scope device
// array.  Every GPU that loads the library holds the K7_NMS_STATIC_BYTES below until exit.  The array's name is only
// a stand-in on the host, so the device address is looked up once per device with cudaGetSymbolAddress (with that
// device current) and kept.  All calls on one device share the array: calls that overlap on the GPU race.
#define K7_NMS_STATIC_BYTES (1u << 20)
static __device__ __align__(256) unsigned char k7_nms_static_scratch[K7_NMS_STATIC_BYTES];

static int k7_nms_static_address(void** out) {
    static std::mutex mu;
    static void* addr[64] = {};
    int dev = -1;
    K7_CUDA(cudaGetDevice(&dev));
    K7_REQUIRE(dev >= 0 && dev < 64, "k7_nms_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        K7_CUDA(cudaGetSymbolAddress(&p, k7_nms_static_scratch));
        K7_REQUIRE(p != nullptr, "k7_nms_static: no device address for the scratch array");
        addr[dev] = p;
    }
    *out = addr[dev];
    return K7_OK;
}

extern "C" int k7_nms_static_max_n(void) { return K7_NMS_STATIC_MAX_N; }

extern "C" int k7_nms_static(const float* boxes, const float* scores, int n, float iou_thr, int64_t* keep, int32_t* count,
                             void* stream) {
    static_assert(K7_NMS_STATIC_BYTES % 256 == 0, "scratch size");
    K7_REQUIRE(n <= K7_NMS_STATIC_MAX_N, "k7_nms_static: n=%d exceeds the built-in scratch (at most %d boxes); use k7_nms",
               n, K7_NMS_STATIC_MAX_N);
    const size_t need = k7_nms_workspace_bytes(K7_NMS_STATIC_MAX_N);
    K7_REQUIRE(need <= K7_NMS_STATIC_BYTES, "k7_nms_static: scratch array of %u bytes is smaller than the %zu needed",
               K7_NMS_STATIC_BYTES, need);
    void* scratch = nullptr;
    int rc = k7_nms_static_address(&scratch);
    if (rc) return rc;
    // k7_nms_run clears the part of the scratch that must start at zero (counters, bounds, bit rows, cell tables).
    return k7_nms_run(boxes, scores, nullptr, n, iou_thr, keep, count, scratch, K7_NMS_STATIC_BYTES, stream);
}

// ------------------------------------------------------------------ Soft-NMS ("soft variants")
// Bodla et al., "Soft-NMS - Improving Object Detection With One Line of Code" (ICCV 2017): instead of
// removing the boxes that overlap the current maximum, their scores decay with the overlap
//     linear:    s_j *= 1 - iou   if iou > thr            gaussian:  s_j *= exp(-iou^2 / sigma)
// and the loop ends when the best remaining score is below score_thr.  Neither the reference nor
// torchvision has it: the restatement in oracle/restate.py follows the paper (parity unpinned, it
// says so).  The picks are inherently sequential (every pick changes the scores the next one is
// taken from), so this is one CTA per call: scores live in shared memory, a pick is a block
// arg-max (ties: lowest index) and a decay pass over the remaining boxes.  Not a bandwidth kernel.
// The overlap is torchvision.ops.nms's: inter / (area_i + area_j - inter), float32, no FMA.
#define K7_SOFT_THREADS 1024
#define K7_SOFT_MAX_N 49152   // scores in shared memory: 192 KB
__global__ void __launch_bounds__(K7_SOFT_THREADS, 1)
k7_soft_nms_kernel(const float4* __restrict__ boxes, const float* __restrict__ scores, int n, int method, float iou_thr,
                   float sigma, float score_thr, int64_t* __restrict__ keep, float* __restrict__ out_scores,
                   int32_t* __restrict__ count) {
    extern __shared__ float s_sc[];
    __shared__ float s_wv[K7_SOFT_THREADS / 32];
    __shared__ int s_wi[K7_SOFT_THREADS / 32];
    __shared__ float s_selv;
    __shared__ int s_sel;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float ninf = -INFINITY;
    for (int j = tid; j < n; j += K7_SOFT_THREADS) {
        const float v = scores[j];
        s_sc[j] = (v == v) ? v : ninf;   // a NaN score is never picked
    }
    __syncthreads();
    int it = 0;
    for (; it < n; ++it) {
        // block arg-max over the remaining scores; ties go to the lowest index
        float bv = ninf;
        int bi = 0x7fffffff;
        for (int j = tid; j < n; j += K7_SOFT_THREADS) {
            const float v = s_sc[j];
            if (v > bv) {
                bv = v;
                bi = j;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_down_sync(0xffffffffu, bv, o);
            const int oi = __shfl_down_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) {
                bv = ov;
                bi = oi;
            }
        }
        if (lane == 0) {
            s_wv[warp] = bv;
            s_wi[warp] = bi;
        }
        __syncthreads();
        if (warp == 0) {
            bv = s_wv[lane];
            bi = s_wi[lane];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_down_sync(0xffffffffu, bv, o);
                const int oi = __shfl_down_sync(0xffffffffu, bi, o);
                if (ov > bv || (ov == bv && oi < bi)) {
                    bv = ov;
                    bi = oi;
                }
            }
            if (lane == 0) {
                s_selv = bv;
                s_sel = bi;
            }
        }
        __syncthreads();
        const int sel = s_sel;
        const float sv = s_selv;
        if (sel == 0x7fffffff || !(sv >= score_thr)) break;   // block-uniform: nothing left worth keeping
        const float4 bi4 = __ldg(boxes + sel);
        const K7Box me = {bi4.x, bi4.y, bi4.z, bi4.w};
        const float area_i = k7_area(bi4.x, bi4.y, bi4.z, bi4.w);
        if (tid == 0) {
            keep[it] = (int64_t)sel;
            out_scores[it] = sv;
            s_sc[sel] = ninf;
        }
        for (int j = tid; j < n; j += K7_SOFT_THREADS) {
            if (j == sel) continue;
            const float v = s_sc[j];
            if (v == ninf) continue;
            const float4 bj = __ldg(boxes + j);
            const float w = fmaxf(0.0f, K7_FSUB(fminf(me.x2, bj.z), fmaxf(me.x1, bj.x)));
            const float h = fmaxf(0.0f, K7_FSUB(fminf(me.y2, bj.w), fmaxf(me.y1, bj.y)));
            const float inter = K7_FMUL(w, h);
            const float iou = K7_FDIV(inter, K7_FSUB(K7_FADD(area_i, k7_area(bj.x, bj.y, bj.z, bj.w)), inter));
            if (!(iou > 0.0f)) continue;   // disjoint or degenerate (NaN): no decay
            float wgt = 1.0f;
            if (method == K7_SOFT_LINEAR) {
                if (iou > iou_thr) wgt = K7_FSUB(1.0f, iou);
            } else {
                wgt = expf(-K7_FDIV(K7_FMUL(iou, iou), sigma));
            }
            s_sc[j] = K7_FMUL(v, wgt);
        }
        __syncthreads();
    }
    if (tid == 0) *count = it;
}

extern "C" int k7_soft_nms(const float* boxes, const float* scores, int n, const K7SoftNmsParams* params, int64_t* keep,
                           float* out_scores, int32_t* count, void* stream) {
    K7_REQUIRE(count && params, "k7_soft_nms: null argument");
    K7_REQUIRE(n >= 0 && n <= K7_SOFT_MAX_N, "k7_soft_nms: n=%d outside [0, %d]", n, K7_SOFT_MAX_N);
    K7_REQUIRE(params->method == K7_SOFT_LINEAR || params->method == K7_SOFT_GAUSSIAN, "k7_soft_nms: bad method");
    K7_REQUIRE(params->method == K7_SOFT_LINEAR || params->sigma > 0.0f, "k7_soft_nms: sigma must be > 0");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (n == 0) {
        K7_CUDA(cudaMemsetAsync(count, 0, sizeof(int32_t), st));
        return K7_OK;
    }
    K7_REQUIRE(boxes && scores && keep && out_scores, "k7_soft_nms: null argument");
    K7_REQUIRE((reinterpret_cast<uintptr_t>(boxes) & 15) == 0, "k7_soft_nms: boxes must be 16-byte aligned");
    const size_t smem = sizeof(float) * (size_t)n;
    K7_CUDA(cudaFuncSetAttribute(k7_soft_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(sizeof(float) * K7_SOFT_MAX_N)));
    k7_soft_nms_kernel<<<1, K7_SOFT_THREADS, smem, st>>>(reinterpret_cast<const float4*>(boxes), scores, n, params->method,
                                                         params->iou_thr, params->sigma, params->score_thr, keep, out_scores,
                                                         count);
    K7_LAUNCH_CHECK();
    return K7_OK;
}
