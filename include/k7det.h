/*
 * k7det.h — C ABI of libk7det.so, the B200 (sm_100a) detection-head hot path.
 *
 * This is the drop-in boundary for the reference's detection-head modules
 * (Kai7/ObjectDetection.Pytorch).  The reference has no FFI of its own — the
 * path is reached through four nn.Module attributes of every detector and one
 * helper function — so each entry point below names the
 * reference callable it replaces; `b200det` (Python) binds them with ctypes
 * and re-creates those modules on top (INTEGRATION.md shows the binding).
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer that is not marked "host"
 *     is a DEVICE pointer on the current CUDA device
 *   - the library never allocates or frees: outputs and workspaces come from
 *     the caller (sizes from the *_workspace_bytes queries)
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*)
 *     and never synchronises the device
 *   - return value: K7_OK or a negative K7_E* code; k7_last_error() gives the
 *     thread-local message of the last failure
 *   - all float tensors are contiguous float32; `cls`/`dcls` base pointers
 *     must be 16-byte aligned (they are streamed with bulk async copies)
 *   - there is no CPU implementation behind this ABI
 *
 * Tensor vocabulary (reference names): B images, A anchors per image, C classes,
 * M annotation rows per image, classifications [B,A,C] (post-sigmoid
 * probabilities, class innermost), regressions [B,A,4], annotations [B,M,5]
 * rows (x1,y1,x2,y2,label) with pad rows label == -1.
 */
#ifndef K7DET_H_
#define K7DET_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define K7_ABI_VERSION 14

#if defined(__GNUC__)
#define K7_API __attribute__((visibility("default")))
#else
#define K7_API
#endif

#define K7_OK 0
#define K7_EINVAL (-1)     /* bad argument (shape, null pointer, alignment, unsupported size) */
#define K7_EWORKSPACE (-2) /* workspace too small */
#define K7_ECUDA (-3)      /* a CUDA runtime call failed */

#define K7_MAX_LEVELS 8
#define K7_MAX_CELL_ANCHORS 16

#define K7_BOX_XYXY 0 /* architectures/anchors.py, losses.py, utils.py */
#define K7_BOX_YXYX 1 /* architectures/utils_efficientdet.py, loss_efficientdet.py */

/* Anchor recipe: everything Anchors.forward (architectures/anchors.py:38-60,
 * generate_anchors :62-97, shift :99-119; yx variant utils_efficientdet.py:74-137)
 * needs to produce anchor `a` without materialising the [A,4] table:
 *   coordinate = float32( base[level][k][c] + (i + 0.5) * stride[level] )      (float64 arithmetic)
 * with anchors ordered level-major, then cell row y, cell column x, then the
 * in-cell index k (ratio-major/scale-minor for xyxy, scale-major/ratio-minor for yxyx). */
typedef struct K7AnchorSpec {
    int32_t num_levels;
    int32_t cell_anchors;                 /* len(ratios) * len(scales) */
    int32_t height, width;                /* image size the recipe was instantiated for */
    int32_t box_order;                    /* K7_BOX_XYXY | K7_BOX_YXYX: layout of base[][][4] and of materialised anchors */
    int32_t total;                        /* A */
    int32_t fh[K7_MAX_LEVELS];            /* ceil(height / stride) (anchors.py:43) */
    int32_t fw[K7_MAX_LEVELS];
    int32_t offset[K7_MAX_LEVELS + 1];    /* first anchor index of each level; offset[num_levels] == A */
    double stride[K7_MAX_LEVELS];
    double base[K7_MAX_LEVELS][K7_MAX_CELL_ANCHORS][4];
} K7AnchorSpec;

/* Every kernel that needs anchors takes BOTH `spec` (host pointer, may be NULL) and
 * `anchors` (device [A,4] float32 in `box_order` layout, may be NULL).  With a spec the
 * anchors are generated in registers and never read from memory; with only a tensor
 * (anchors built by someone else) they are loaded.  One of the two must be given. */

K7_API int k7_version(void);
K7_API const char* k7_last_error(void);
/* Kernels this library has launched in this process so far (all entry points, all threads): what a caller
 * subtracts around a region to count launches.  Memsets and copies are not kernels and not counted.  No
 * counterpart in the reference (measurement aid, bench.py's gpu_launches). */
K7_API unsigned long long k7_launch_count(void);

/* Replaces Anchors.forward (architectures/anchors.py:38-60; utils_efficientdet.py:74-137):
 * writes out[A,4] float32. */
K7_API int k7_anchors_materialize(const K7AnchorSpec* spec, float* out, void* stream);

/* ---- training side ------------------------------------------------------------------ */

typedef struct K7MatchParams {
    float neg_thr; /* IoU_max <  neg_thr -> negative   (losses.py:79: 0.35; loss_efficientdet.py:75: 0.4) */
    float pos_thr; /* IoU_max >= pos_thr -> positive   (losses.py:80: 0.45; loss_efficientdet.py:77: 0.5) */
} K7MatchParams;

#define K7_STATE_NEG 0
#define K7_STATE_IGNORE 1
#define K7_STATE_POS 2
/* assign word: bits 0-1 = state, bits 2-31 = annotation row index (0..M-1) of the arg-max GT.
 * The row is the reference's IoU_argmax for POSITIVE anchors (the only ones the reference uses it
 * for, losses.py:94-97, 125); for the others it is exact only when iou_max/argmax are requested
 * (without them the kernel skips ground-truth boxes that provably stay below neg_thr). */
#define K7_ASSIGN_STATE(w) ((w) & 3)
#define K7_ASSIGN_ROW(w) ((int32_t)((uint32_t)(w) >> 2))

/* Replaces calc_iou + IoU-max assignment (losses.py:6-23, 49-82).  Outputs:
 *   assign  [B,A] int32   packed state / GT row (images without a valid row: all IGNORE).
 *   num_pos [B]   int32   positives per image
 *   iou_max [B,A] float32 (optional, may be NULL)  bit-identical to torch.max(IoU,1).values
 *   argmax  [B,A] int32   (optional, may be NULL)  index into the FILTERED annotation list
 *   zero_fill4 [B,A,4] float32 (optional, may be NULL) is filled with zeros: the d(reg) buffer a
 *                         following k7_loss will write positive rows into (the kernel is ALU-bound,
 *                         so the 16 B/anchor store is free here; see K7LossParams.dreg_prezeroed)
 *   ignore_bits [B, K7_IGNORE_WORDS(A)] uint32 (optional, may be NULL; 16-byte aligned) one bit per
 *                         anchor, set = state IGNORE: bit a%32 of word a/32 of the image's row (a compact
 *                         form of the states for callers that want it; k7_loss does not need it). */
#define K7_IGNORE_WORDS(A) (((((size_t)(A)) + 127) >> 7) << 2) /* words per image, a multiple of 16 bytes */
K7_API int k7_match(const K7AnchorSpec* spec, const float* anchors, int box_order,
             const float* annotations, int B, int A, int M, const K7MatchParams* params,
             int32_t* assign, int32_t* num_pos, float* iou_max, int32_t* argmax, float* zero_fill4,
             uint32_t* ignore_bits, void* stream);

/* The same assignment computed GROUND-TRUTH-FIRST - the production path of FocalLoss when the anchors come
 * from a recipe (calc_iou + IoU-max assignment, losses.py:6-23, 49-82).  An anchor only leaves the NEGATIVE
 * state through a ground-truth box with IoU >= neg_thr, and for one box those anchors are a few cells
 * around it on the two or three pyramid levels whose anchor areas are within 1/neg_thr of its own.  One warp
 * per (image, annotation row) enumerates a provable superset of them (area window and the intersection
 * width / height any pair with IoU >= neg_thr needs, 1e-4 margins; all shapes' candidate cells as one flat
 * range over the lanes), evaluates the reference's float32 IoU (k7_pair_iou: same operations, same order)
 * and folds pairs with IoU >= neg_thr into
 *     best[b*A + a] = max( generation << 48 | IoU bits << 16 | (0xFFFF - annotation row) )        (atomicMax)
 * i.e. the largest IoU and, on ties, the lowest row - torch.max's first index (losses.py:66).  Work is
 * proportional to the pairs that matter, not to A x M, and nothing per anchor is written: an entry whose
 * top 16 bits differ from `generation` means "no ground truth reached neg_thr" = NEGATIVE.
 *   best        caller-owned buffer of k7_match_scatter_workspace_bytes(B, A) bytes, 256-byte aligned,
 *               ZERO-INITIALISED ONCE (and again whenever B or A change); every call on it must pass a LARGER
 *               generation than the one before (1 .. K7_BEST_GEN_MAX; re-zero before starting over).  Layout:
 *               [B*A] uint64 maxima, then - at K7_BEST_LIST_OFFSET(B, A) - a header with one uint32 counter per
 *               image, K7_BEST_COUNT_STRIDE words apart (atomics on one address serialise; one counter per
 *               image keeps them parallel), then - K7_BEST_LIST_HEADER(B) bytes further - B lists of up to A
 *               uint32 anchor indices: list b holds the anchors of image b whose entry was written for the
 *               first time in this generation (the value atomicMax returns tells), in no particular order.
 *               k7_loss reads the lists (K7LossParams.match_generation / neg_thr / pos_thr) and through them
 *               only the few per cent of entries that matter; nothing walks the B*A array.
 *   zero_fill4  [B,A,4] float32 (optional, may be NULL; 16-byte aligned) is filled with zeros: the d(reg)
 *               buffer of the k7_loss call that follows (K7LossParams.dreg_prezeroed).  The kernel is
 *               latency-bound, idle warps do the stores.
 * Requires a recipe, 1e-3 < neg_thr <= pos_thr, M <= 65535, B <= 65535.  Images without a valid row leave
 * `best` untouched; k7_loss sees from the annotations that the image is skipped (losses.py:52-60). */
#define K7_BEST_GEN_MAX 65535u
#define K7_BEST_GEN(v) ((uint32_t)((v) >> 48))
#define K7_BEST_IOU_BITS(v) ((uint32_t)((v) >> 16))
#define K7_BEST_ROW(v) (0xFFFF - (int32_t)((v) & 0xFFFFu))
#define K7_BEST_LIST_OFFSET(B, A) ((((size_t)8 * (size_t)(B) * (size_t)(A)) + 255) / 256 * 256)
#define K7_BEST_COUNT_STRIDE 16
#define K7_BEST_LIST_HEADER(B) ((((size_t)4 * K7_BEST_COUNT_STRIDE * (size_t)(B)) + 255) / 256 * 256)
K7_API size_t k7_match_scatter_workspace_bytes(int B, int A);
K7_API int k7_match_scatter(const K7AnchorSpec* spec, const float* annotations, int B, int A, int M,
                     const K7MatchParams* params, uint64_t* best, uint32_t generation, float* zero_fill4,
                     void* stream);

/* Multi-GPU: the path's one exchange - the four batch sums {sum cls_i, sum reg_i, sum num_pos_i, B}
 * of every rank - fused into k7_loss.  mailbox[p] is a device pointer, valid in THIS process, to
 * rank p's mailbox (peer memory over NVLink: cudaIpcOpenMemHandle / cuMem import; mailbox[rank] is
 * the local one): 2 * world slots of 8 floats, zero-initialised before the first call.  The last
 * block of the row kernel posts this rank's sums into every mailbox and adds the world posts it
 * receives in rank order, so `sums` holds the GLOBAL sums, bit-identical on every rank, with no
 * separate all-reduce launch.  seq counts the calls (1, 2, ...) and must be the same on every rank (or give
 * seq_counter and let the kernel count).
 * A peer that does not post within timeout_ms (default 10 000) gives NaN sums, never a hang. */
#define K7_MAX_PEERS 16
typedef struct K7PeerReduce {
    int32_t world, rank;
    uint32_t seq;
    int32_t timeout_ms;
    float* mailbox[K7_MAX_PEERS];
    uint32_t* seq_counter; /* optional: TWO device words (4-byte aligned, zero before the first call) - calls posted, calls
                            * collected.  When given, the call number is taken from them and advanced by the kernels,
                            * `seq` is ignored - a CUDA graph that captured the call can then be replayed (a kernel
                            * argument would be frozen into the graph) */
    int32_t post_only;     /* 1: k7_loss only POSTS this rank's sums into the mailboxes and writes its OWN sums to `sums`;
                            * nothing in the call waits for a peer.  The global sums are fetched later, off the critical
                            * path, with k7_peer_collect (the gradients never need them: they are pre-scaled with the
                            * statically known global batch).  Needs seq_counter; the word behind it (seq_counter[1])
                            * counts the collects.  Order: collect k must be enqueued before k7_loss call k+1 on streams
                            * that order them (two slot sets alternate: a rank must not post call k+2 before it has
                            * collected call k+1). */
} K7PeerReduce;
#define K7_PEER_MAILBOX_FLOATS(world) (2 * (world) * 8)
#define K7_IPC_HANDLE_BYTES 64

/* Mailbox plumbing (CUDA IPC; one process per GPU on one node).  create: a zeroed mailbox on the
 * current device + its 64-byte handle, which the host sends to the other ranks by any transport;
 * open: maps a peer's mailbox into this process and enables peer access from the current device;
 * close / destroy undo open / create.  Not stream-ordered: set-up time only. */
K7_API int k7_peer_mailbox_create(int world, float** mailbox, unsigned char* handle /* [K7_IPC_HANDLE_BYTES] */);
K7_API int k7_peer_mailbox_open(const unsigned char* handle, float** mailbox);
K7_API int k7_peer_mailbox_close(float* mailbox);
K7_API int k7_peer_mailbox_destroy(float* mailbox);
/* The second half of a post_only exchange: one small kernel (W threads) that waits - on `stream`, which may be any
 * stream - until every rank's post of the next uncollected call has arrived in this rank's mailbox and writes the
 * four GLOBAL sums, added in rank order (the same bits on every rank), to sums[4].  NaN after timeout_ms.
 * Replaces the gather + .mean() of nn.DataParallel (train_ksevendet.py:358-361) like K7PeerReduce itself. */
K7_API int k7_peer_collect(const K7PeerReduce* reduce, float* sums, void* stream);

typedef struct K7LossParams {
    float alpha;            /* 0.25 (losses.py:31) ; gamma is fixed at 2 (losses.py:32, pow == x*x) */
    float beta;             /* smooth-L1 switch 1/9 (losses.py:157-159) */
    float target_div[4];    /* targets / [0.1,0.1,0.2,0.2] (losses.py:148); ones for loss_efficientdet.py */
    int32_t yx_targets;     /* 0: (dx,dy,dw,dh) (losses.py:144)  1: (dy,dx,dh,dw) (loss_efficientdet.py:127) */
    int32_t mean_divisor;   /* images the batch mean is taken over: B for one GPU, the GLOBAL batch when sharded */
    int32_t dreg_prezeroed; /* 1: the caller (k7_match's zero_fill4) already zeroed dreg; 0: k7_loss zeroes it itself */
    int32_t from_logits;    /* 0: cls holds post-sigmoid probabilities (the reference's tensor).  1: cls holds the
                             * head's pre-sigmoid logits (output_act of retinanet.py:140,159 / header_act of
                             * simple_head.py:310 removed): the sigmoid is fused into the stream and dcls is
                             * d(loss)/d(logit).  SURVEY 8(f) row 1. */
    const K7PeerReduce* peer_reduce; /* NULL: `sums` are this rank's own; else see K7PeerReduce */
    uint32_t match_generation; /* `best` given: the generation of the k7_match_scatter call that filled it */
    float neg_thr, pos_thr;    /* `best` given: the thresholds of that call (the states are decided from the IoU bits) */
    int32_t* assign_out;       /* optional [B,A] int32: the packed state / row words of k7_match, written on the way */
    int32_t spare_sms;         /* the stream kernel (one persistent CTA per SM) leaves this many SMs unused: a caller that
                                * overlaps the call with SM-bound work of another stream (the NMS kernels of a k7_detect
                                * on few images: their 1024-thread CTAs cannot share an SM with a stream CTA) gives them
                                * room at no cost - up to ~16 of 148 SMs the stream stays HBM-bound.  0: all SMs. */
    int32_t consume_best;      /* 1: entries of this generation are zeroed as they are read, so the SAME generation can be
                                * used again by the next k7_match_scatter on this array (a captured CUDA graph replays
                                * one generation); `best` is then written by k7_loss */
} K7LossParams;

K7_API size_t k7_loss_workspace_bytes(int B, int A, int C);

/* Replaces FocalLoss.forward after assignment (losses.py:62, 94-168) and, when dcls/dreg
 * are given, its autograd backward.  The assignment comes either as k7_match's words (`assign`) or as
 * k7_match_scatter's `best` buffer (exactly one of the two is non-NULL).  Three kernels: a list kernel that turns
 * the touched lists of `best` (or, for `assign`, one pass over the words) into a queue of positive rows and a
 * one-bit-per-anchor bitmap of ignored rows (neg_thr <= IoU_max < pos_thr, losses.py:79-82); a stream over `cls`
 * that evaluates every element as "target 0" at HBM speed and skips the ignored rows by the bitmap; a serve kernel
 * that gives the positive rows smooth-L1 + d(reg) and moves their label element from target 0 to target 1.  All
 * contributions are accumulated as 2^-32 fixed-point integers per image (order independent): results are
 * bit-reproducible run to run.  Outputs:
 *   num_pos   [B]   int32    positives per image, counted while the assignment is walked (it is WRITTEN here,
 *                            whichever form the assignment comes in)
 *   per_image [B,2] float32  (cls_loss_i, reg_loss_i)
 *   sums      [4]   float32  { sum_i cls_loss_i, sum_i reg_loss_i, sum_i num_pos_i, B }  (the NCCL payload)
 *   dcls [B,A,C], dreg [B,A,4] (optional, both or neither): d(mean cls loss)/dcls and
 *   d(mean reg loss)/dreg with the mean taken over params->mean_divisor images.
 * The batch means of losses.py:168 are sums[0]/B and sums[1]/B.
 * Any B, A, C is streamed with bulk copies; only the base pointers of cls / dcls / reg / dreg must be
 * 16-byte aligned. */
K7_API int k7_loss(const K7AnchorSpec* spec, const float* anchors, int box_order,
            const float* cls, const float* reg, const float* annotations,
            const int32_t* assign, const uint64_t* best, int32_t* num_pos,
            int B, int A, int C, int M, const K7LossParams* params,
            float* per_image, float* sums, float* dcls, float* dreg,
            void* workspace, size_t workspace_bytes, void* stream);

/* The same on the classification head's RAW outputs: one NCHW tensor per pyramid level,
 * levels[l] = [B, cell_anchors*C, fh[l], fw[l]] float32 (channel = in-cell anchor * C + class), i.e. what the
 * head's last convolution writes BEFORE permute(0,2,3,1) / view / cat (retinanet.py:162-178, 286;
 * simple_head.py:289-307) - those passes, and with params->from_logits the sigmoid, never touch memory.
 * `levels` / `dlevels` are HOST arrays of spec->num_levels device pointers (16-byte aligned); dlevels[l]
 * (optional, with dreg) receives d(loss)/d(levels[l]) in the same layout.  An anchor spec is required
 * (the level geometry comes from it); regressions stay [B,A,4]. */
K7_API size_t k7_loss_levels_workspace_bytes(const K7AnchorSpec* spec, int B, int C);
K7_API int k7_loss_levels(const K7AnchorSpec* spec, const float* const* levels, const float* reg,
                   const float* annotations, const int32_t* assign, const uint64_t* best,
                   int32_t* num_pos, int B, int C, int M, const K7LossParams* params,
                   float* per_image, float* sums, float* const* dlevels, float* dreg,
                   void* workspace, size_t workspace_bytes, void* stream);

/* grad *= *scale (device scalar); no-op when *scale == 1.  Used by the autograd backward
 * when the incoming gradient of a loss output is not 1. */
K7_API int k7_scale_inplace(float* grad, size_t n, const float* scale, void* stream);

/* The same for up to K7_SCALE_MAX tensors in one launch: grads[i][0..n[i]) *= *scales[i] (host arrays of
 * device pointers / sizes).  The autograd backward of a step scales d(classifications) and d(regressions)
 * - or the per-level gradients - with it. */
#define K7_SCALE_MAX 8
K7_API int k7_scale_inplace_multi(float* const* grads, const size_t* n, const float* const* scales, int count,
                           void* stream);

/* ---- inference side ----------------------------------------------------------------- */

#define K7_DECODE_RETINA 0 /* utils.py:26-53: (dx,dy,dw,dh) * std + mean around xyxy anchors */
#define K7_DECODE_EFFDET 1 /* utils_efficientdet.py:8-35: (dy,dx,dh,dw), no mean/std, yxyx anchors */

typedef struct K7DecodeParams {
    int32_t variant;
    float mean[4], std[4];
    int32_t clip;           /* fuse ClipBoxes (utils.py:61-71) */
    int32_t height, width;
} K7DecodeParams;

/* Replaces BBoxTransform.forward (+ optionally ClipBoxes.forward): boxes[B,A,4] (x1,y1,x2,y2).
 * anchor_batch_stride: 0 when `anchors` is [1,A,4] (broadcast), A*4 when it is [B,A,4]. */
K7_API int k7_decode(const K7AnchorSpec* spec, const float* anchors, int box_order, size_t anchor_batch_stride,
              const float* deltas, int B, int A, const K7DecodeParams* params, float* boxes, void* stream);

/* Replaces ClipBoxes.forward (utils.py:61-71), in place. */
K7_API int k7_clip(float* boxes, int B, int A, int height, int width, void* stream);

typedef struct K7DetectParams {
    K7DecodeParams decode;  /* clip/height/width are honoured */
    float score_thr;        /* strict '>' (retinanet.py:302: 0.05) */
    float iou_thr;          /* strict '>' (torchvision.ops.nms, retinanet.py:314: 0.5) */
    int32_t pre_nms_topk;   /* 0 = off (reference behaviour) */
    int32_t max_detections; /* 0 = off; otherwise stop after this many kept boxes per image */
    int32_t from_logits;    /* 1: cls holds pre-sigmoid logits; scores are 1/(1+expf(-z)) (torch.sigmoid), the
                             * threshold and the arg-max ties are decided on those float32 probabilities */
    int32_t class_aware;    /* 0: class-agnostic NMS, the reference (retinanet.py:314); 1: a box is suppressed only by
                             * a higher-scored box of the SAME arg-max class (torchvision.ops.batched_nms with
                             * idxs = labels, per-class strategy) - SURVEY 8f row 4 */
    int32_t nms_chain;      /* which NMS kernels run (same result): 0 = automatic - the per-image chain (one CTA per image
                             * sorts and resolves, k7_nms_v1.cu) when the batch has >= 12 images or images are small
                             * (A < 262 144), else the multi-SM chain (every image spread over all SMs: a few images
                             * with thousands of survivors each, a shard of a large batch); 1 = per-image, 2 = multi-SM */
} K7DetectParams;
#define K7_NMS_CHAIN_AUTO 0
#define K7_NMS_CHAIN_PER_IMAGE 1
#define K7_NMS_CHAIN_MULTI_SM 2

K7_API size_t k7_detect_workspace_bytes(int B, int A, int C);

/* Replaces the eval tail of the detectors for EVERY image of the batch
 * (retinanet.py:297-318; batched form utils/utils.py:76-114): fused decode + clip +
 * max-over-class + score threshold + compaction, then per image a radix sort
 * (score descending, anchor ascending) and greedy bitmask NMS.  Results stay in the
 * workspace; `counts` gets {survivors, kept} per image:  counts[B,2] int32. */
K7_API int k7_detect(const K7AnchorSpec* spec, const float* anchors, int box_order,
              const float* cls, const float* reg, int B, int A, int C,
              const K7DetectParams* params, int32_t* counts,
              void* workspace, size_t workspace_bytes, void* stream);

/* k7_detect on the head's per-level NCHW outputs (see k7_loss_levels); same workspace, same results. */
K7_API int k7_detect_levels(const K7AnchorSpec* spec, const float* const* levels, const float* reg, int B, int C,
                     const K7DetectParams* params, int32_t* counts, void* workspace, size_t workspace_bytes,
                     void* stream);

/* Copies the kept detections of image b to row-major outputs with `cap` rows per image:
 *   scores [B,cap] f32, labels [B,cap] i64, boxes [B,cap,4] f32,
 *   anchor_ids [B,cap] i32 (optional), survivor_rank [B,cap] i32 (optional: the index
 *   torchvision.ops.nms would have returned, i.e. position in the thresholded list).
 * Rows beyond the image's kept count are left untouched; kept boxes beyond `cap` are dropped. */
K7_API int k7_detect_gather(const void* workspace, size_t workspace_bytes, int B, int A, int C, int cap,
                     float* scores, int64_t* labels, float* boxes, int32_t* anchor_ids,
                     int32_t* survivor_rank, void* stream);

/* Stand-alone replacement of torchvision.ops.nms(boxes, scores, thr) for one image
 * (call sites retinanet.py:314 etc.): keep[n] int64 indices in descending-score order,
 * count[1] int32.  Workspace: k7_nms_workspace_bytes(n). */
K7_API size_t k7_nms_workspace_bytes(int n);
K7_API int k7_nms(const float* boxes, const float* scores, int n, float iou_thr,
           int64_t* keep, int32_t* count, void* workspace, size_t workspace_bytes, void* stream);

/* Per-class NMS for one image (the reference itself is class-agnostic): the result of
 * torchvision.ops.batched_nms(boxes, scores, idxs, thr) computed per category ("vanilla" strategy,
 * torchvision/ops/boxes.py _batched_nms_vanilla): box j is suppressed iff a kept box i of the same idxs value
 * with a higher score (ties: lower index) has IoU(i, j) > thr.  keep[n] int64 indices in descending-score order
 * (ties: lower index first), count[1] int32.  idxs [n] int64 (values must fit 32 bits).  Same kernels and
 * workspace as k7_nms: the category test sits after the division-free filter of the bitmask kernel. */
K7_API int k7_batched_nms(const float* boxes, const float* scores, const int64_t* idxs, int n, float iou_thr,
                   int64_t* keep, int32_t* count, void* workspace, size_t workspace_bytes, void* stream);

/* k7_nms for callers that may not allocate device memory (code under CUDA graph capture, serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.
 * Same inputs, outputs and results as k7_nms, but the scratch is a fixed device array inside the library
 * (1 MiB, held by every GPU that loads the library until the process exits), so the entry point takes no
 * workspace and calls no cudaMalloc / cudaFree / allocating routine.  Limit: n <= K7_NMS_STATIC_MAX_N
 * (= k7_nms_static_max_n()); a larger n returns K7_EINVAL with a message naming the limit.  All calls on one
 * device share the array: it is cleared where it must start at zero by every call, and two calls that overlap
 * on the GPU (different streams, no ordering between them) race - serialise them, or use k7_nms with one
 * workspace per stream. */
#define K7_NMS_STATIC_MAX_N 4096
K7_API int k7_nms_static_max_n(void);
K7_API int k7_nms_static(const float* boxes, const float* scores, int n, float iou_thr,
                  int64_t* keep, int32_t* count, void* stream);

/* Soft-NMS for one image (Bodla et al., ICCV 2017; SURVEY 8f row 4 "soft variants" - neither the reference nor
 * torchvision has one, so the behaviour is the paper's, restated in oracle/restate.py::soft_nms): repeatedly pick
 * the box with the highest remaining score (ties: lowest index) and decay the scores of the others by their overlap
 * with it (torchvision.ops.nms's IoU),
 *     K7_SOFT_LINEAR:   s_j *= 1 - iou  if iou > iou_thr        K7_SOFT_GAUSSIAN:  s_j *= exp(-iou^2 / sigma)
 * until the best remaining score is below score_thr.  keep[n] int64 indices in pick order, out_scores[n] float32 the
 * score each box had when it was picked, count[1] int32.  n <= 49 152 (the scores live in shared memory).  One CTA:
 * the picks are sequential by definition; this is an extension for callers who want it, not part of the hot path. */
#define K7_SOFT_LINEAR 0
#define K7_SOFT_GAUSSIAN 1
typedef struct K7SoftNmsParams {
    int32_t method;    /* K7_SOFT_LINEAR | K7_SOFT_GAUSSIAN */
    float iou_thr;     /* linear: overlaps above it decay (0.3 in the paper) */
    float sigma;       /* gaussian (0.5 in the paper) */
    float score_thr;   /* stop when the best remaining score is below it (0.001 in the paper) */
} K7SoftNmsParams;
K7_API int k7_soft_nms(const float* boxes, const float* scores, int n, const K7SoftNmsParams* params,
                int64_t* keep, float* out_scores, int32_t* count, void* stream);

/* ---------------------------------------------------------------- annotation packing
 * collater's annotation half (datasettool/dataloader.py:678-686): rows [N,5] of all images back
 * to back (device), offsets [B+1] int32 (device) -> out [B,M,5], rows beyond an image's count are
 * -1 in every field.  M = max(max_b n_b, 1) as the reference (:676, :688); an image with more than
 * M rows is truncated.  One launch; the host sends one ragged block instead of padding in Python. */
K7_API int k7_annotations_pad(const float* rows, const int32_t* offsets, int B, int M, float* out, void* stream);

/* ---------------------------------------------------------------- evaluator tails on device
 * What the reference's evaluators do on the host with every image's (scores, labels, boxes)
 * after a device->host copy, for a whole batch in three small launches:
 *   K7_EVAL_COCO  datasettool/coco_eval.py:25-57, ksevendata_eval.py:24-61
 *       boxes / scale, xyxy -> xywh (float32), labels outside `valid_label` skipped, rows of an
 *       image stop at the first float(score) < score_thr (float64 comparison, as Python does);
 *       row = (image_id, x, y, w, h, score, category_of_label[label]), images in order.
 *   K7_EVAL_CSV   datasettool/csv_eval.py:93-121 (_get_detections)
 *       boxes / scale, rows with score > (float)score_thr, the first max_detections of them
 *       (input order = descending score), bucketed by label:
 *       row = (image index, x1, y1, x2, y2, score, label), ordered by (image, label, rank);
 *       class_offsets[b*num_classes + c] .. [+1] delimit all_detections[b][c].
 * Rows are float64 [N,7] (the float32 values widened exactly: what json.dump / np.concatenate
 * see in the reference); row_offsets[B+1] int32 are the exclusive per-image offsets, N = row_offsets[B].
 * Inputs are the row-padded outputs of k7_detect_gather (or of any detector): scores [B,cap] f32
 * descending per image, labels [B,cap] i64, boxes [B,cap,4] f32, counts[b*counts_stride] rows valid.
 * `scale` holds float32(scale) per image.  Labels outside [0, num_classes) are dropped. */
#define K7_EVAL_COCO 0
#define K7_EVAL_CSV 1

typedef struct K7EvalParams {
    int32_t mode;           /* K7_EVAL_COCO | K7_EVAL_CSV */
    int32_t max_detections; /* CSV: csv_eval.py:64 default 100; 0 = no cap.  COCO: ignored */
    int32_t num_classes;
    int32_t reserved;
    double score_thr;       /* coco_eval.py:6 / csv_eval.py:64 default 0.05 */
} K7EvalParams;

K7_API size_t k7_eval_pack_workspace_bytes(int B, int num_classes);
K7_API int k7_eval_pack(const float* scores, const int64_t* labels, const float* boxes, int cap,
                 const int32_t* counts, int counts_stride, int B, const float* scale,
                 const int64_t* image_ids /* NULL: the image index */,
                 const int32_t* category_of_label /* [num_classes] or NULL: the label */,
                 const uint8_t* valid_label /* [num_classes] or NULL: all (COCO mode) */,
                 const K7EvalParams* params, double* rows /* [B*cap,7] */, int32_t* row_offsets /* [B+1] */,
                 int32_t* class_offsets /* CSV: [B*num_classes+1]; else NULL */,
                 void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* K7DET_H_ */
