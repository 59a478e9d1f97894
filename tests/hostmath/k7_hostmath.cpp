// TEST-ONLY host build of csrc/k7_math.cuh: lets the CPU test-suite check the scalar
// arithmetic the CUDA kernels use (IoU bits, focal polynomial, smooth-L1, decode, NMS
// predicate) against the oracle without a GPU.  Not linked into libk7det.so, not a fallback.
// Build: g++ -O2 -ffp-contract=off -shared -fPIC (tests/test_hostmath.py does it).
#include "../../objectdetection.pytorch_b200/csrc/k7_math.cuh"

extern "C" {

// Per-anchor assignment exactly as k7_match_kernel does it (without the warp-hull culling,
// which only skips pairs whose IoU is 0).  gt = nv valid rows (x1,y1,x2,y2).
void hm_match(const float* anchors, int A, const float* gt, int nv, float neg_thr, float pos_thr,
              float* iou_max, int* argmax, int* state) {
    for (int a = 0; a < A; ++a) {
        const float ax1 = anchors[4 * a], ay1 = anchors[4 * a + 1], ax2 = anchors[4 * a + 2], ay2 = anchors[4 * a + 3];
        const float area_a = k7_area(ax1, ay1, ax2, ay2);
        float best = 0.0f;
        int bi = 0;
        for (int j = 0; j < nv; ++j) {
            const float* g = gt + 4 * j;
            const float iou = k7_pair_iou(ax1, ay1, ax2, ay2, area_a, g[0], g[1], g[2], g[3], k7_area(g[0], g[1], g[2], g[3]));
            if (iou > best) { best = iou; bi = j; }
        }
        int st = 1;
        if (best < neg_thr) st = 0;
        if (best >= pos_thr) st = 2;
        iou_max[a] = best; argmax[a] = bi; state[a] = st;
    }
}

// focal terms of n probabilities treated as target 0 / target 1: loss and d loss/d x
void hm_focal(const float* x, int n, float alpha, float* neg, float* neg_grad, float* pos, float* pos_grad) {
    const float oma = 1.0f - alpha;
    for (int i = 0; i < n; ++i) {
        const float p = k7_clamp_p(x[i]);
        const float q = 1.0f - p;
        const float nl = k7_neg_log1m(q);
        const float pass = k7_clamp_pass(x[i]) ? 1.0f : 0.0f;
        neg[i] = k7_focal_neg(p, nl, oma);
        neg_grad[i] = pass * k7_focal_neg_grad(p, q, nl, oma);
        pos[i] = k7_focal_pos(p, alpha);
        pos_grad[i] = pass * k7_focal_pos_grad(p, alpha);
    }
}

void hm_reg(const float* anchor, const float* gt, const float* reg, const float* div, int yx, float beta, float quad,
            float shift, float* targets, float* loss, float* grad) {
    K7Reg4 t = k7_reg_targets(anchor[0], anchor[1], anchor[2], anchor[3], gt[0], gt[1], gt[2], gt[3], div, yx);
    for (int i = 0; i < 4; ++i) {
        targets[i] = t.v[i];
        loss[i] = k7_smooth_l1(t.v[i], reg[i], beta, quad, shift, grad + i);
    }
}

void hm_decode(const float* anchors_xyxy, const float* deltas, int n, int variant, const float* mean, const float* std,
               int clip, int height, int width, float* out) {
    for (int i = 0; i < n; ++i) {
        const float* a = anchors_xyxy + 4 * i;
        const float* d = deltas + 4 * i;
        K7Box b = variant == 0 ? k7_decode_retina(a[0], a[1], a[2], a[3], d[0], d[1], d[2], d[3], mean, std)
                               : k7_decode_effdet(a[0], a[1], a[2], a[3], d[0], d[1], d[2], d[3]);
        if (clip) b = k7_clip_box(b, height, width);
        out[4 * i] = b.x1; out[4 * i + 1] = b.y1; out[4 * i + 2] = b.x2; out[4 * i + 3] = b.y2;
    }
}

// greedy NMS over boxes visited in the order of the kernel's sort key
int hm_nms(const float* boxes, const float* scores, int n, float thr, long long* keep, unsigned long long* keys_out) {
    // selection sort by key is fine for test sizes
    unsigned long long* keys = keys_out;
    for (int i = 0; i < n; ++i) keys[i] = ((unsigned long long)k7_score_key_desc(scores[i]) << 32) | (unsigned)i;
    for (int i = 1; i < n; ++i) {   // insertion sort
        unsigned long long k = keys[i];
        int j = i - 1;
        while (j >= 0 && keys[j] > k) { keys[j + 1] = keys[j]; --j; }
        keys[j + 1] = k;
    }
    int kept = 0;
    for (int p = 0; p < n; ++p) {
        const int i = (int)(keys[p] & 0xffffffffu);
        const K7Box bi = {boxes[4 * i], boxes[4 * i + 1], boxes[4 * i + 2], boxes[4 * i + 3]};
        const float ai = k7_area(bi.x1, bi.y1, bi.x2, bi.y2);
        bool dead = false;
        for (int k = 0; k < kept && !dead; ++k) {
            const int j = (int)keep[k];
            const K7Box bj = {boxes[4 * j], boxes[4 * j + 1], boxes[4 * j + 2], boxes[4 * j + 3]};
            dead = k7_nms_suppress(bj, k7_area(bj.x1, bj.y1, bj.x2, bj.y2), bi, ai, thr);
        }
        if (!dead) keep[kept++] = i;
    }
    return kept;
}
}
