// Interface of the per-image NMS chain (k7_nms_v1.cu) towards k7_detect.cu.
#pragma once
#include "k7_common.cuh"

#define K7_V1_MIN_IMAGES 12        // automatic choice: batches of at least this many images take the per-image chain,
#define K7_V1_SMALL_IMAGE 262144   // and so do images with fewer anchors than this (few survivors per image)

struct K7V1Ws {             // its workspace areas (behind the areas of k7_detect.cu)
    float4* sbox;           // [B][cap] boxes in spatial order
    int32_t* srank;         // [B][cap] rank (position in score order) of the box at each spatial index
    int32_t* slabel;        // [B][cap] class of the box at each spatial index (class-aware NMS only)
    float4* blk;            // [B][cap/32][2] summaries per 32 boxes
    int32_t* work;          // [B] work-item counter of the mask kernel
    unsigned long long* nz; // [B][cap][nzw] which words of a bitmask row were written
    unsigned long long* M;  // [B][cap][cap/64] suppression bitmask, sparsely written
    size_t bytes;
    int cap;                // survivors per image the chain is sized for; images above it: chunked fallback
};

namespace k7v1 {
K7V1Ws workspace(void* base, int B, int A);
// sort -> mask -> resolve on the survivors the score kernel left: keys0[b][0 .. surv_count[b]) in arrival order,
// box / label by anchor id.  Writes keys1 (rank order), nsel, kept_key / kept_box / kept_count and counts [B,2].
int chain(int B, int A, float iou_thr, int pre_topk, int max_det, bool class_aware, int32_t* surv_count, uint64_t* keys0,
          uint64_t* keys1, const float4* box, const int32_t* label, int32_t* nsel, uint64_t* kept_key, float4* kept_box,
          int32_t* kept_count, int32_t* counts, const K7V1Ws& v, cudaStream_t st);
}  // namespace k7v1
