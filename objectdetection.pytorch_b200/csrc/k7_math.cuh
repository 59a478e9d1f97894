// Scalar arithmetic of the hot path, written once for device and host.
//
// Everything whose RESULT BITS decide an integer output (assignment, arg-max, survivor
// set, NMS keep) is evaluated with explicitly rounded single operations in the order of
// the reference's separate ATen ops, never contracted into FMAs:
//   IoU                  architectures/losses.py:6-23 (loss_efficientdet.py:10-24)
//   regression targets   architectures/losses.py:125-148
//   box decode / clip    architectures/utils.py:26-71, utils_efficientdet.py:8-50
//   NMS overlap test     torchvision.ops.nms (inter / (area_i + area_j - inter) > thr)
// Loss values (float outputs, 1e-5 relative contract) may use FMAs and polynomials.
//
// The host build of this header (tests/hostmath) exists so the arithmetic can be checked
// against the oracle on a machine without a GPU; it is not a product code path.
#pragma once

#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define K7_HD __host__ __device__ __forceinline__
#else
#define K7_HD inline
#endif

#if defined(__CUDA_ARCH__)
#define K7_FADD(a, b) __fadd_rn((a), (b))
#define K7_FSUB(a, b) __fsub_rn((a), (b))
#define K7_FMUL(a, b) __fmul_rn((a), (b))
#define K7_FDIV(a, b) __fdiv_rn((a), (b))
#else
// host: compile with -ffp-contract=off so these stay separate IEEE operations
#define K7_FADD(a, b) ((float)(a) + (float)(b))
#define K7_FSUB(a, b) ((float)(a) - (float)(b))
#define K7_FMUL(a, b) ((float)(a) * (float)(b))
#define K7_FDIV(a, b) ((float)(a) / (float)(b))
#endif

// ------------------------------------------------------------------ IoU (training side)
// area = (x2-x1)*(y2-y1)                                                    losses.py:7,16
K7_HD float k7_area(float x1, float y1, float x2, float y2) {
    return K7_FMUL(K7_FSUB(x2, x1), K7_FSUB(y2, y1));
}

// IoU of anchor a and ground truth g.  When the clamped intersection is empty the
// reference produces exactly +0 (0 / max(ua,1e-8)), so the division is skipped.
K7_HD float k7_pair_iou(float ax1, float ay1, float ax2, float ay2, float area_a, float gx1, float gy1,
                        float gx2, float gy2, float area_g) {
    const float iw = K7_FSUB(fminf(ax2, gx2), fmaxf(ax1, gx1));
    const float ih = K7_FSUB(fminf(ay2, gy2), fmaxf(ay1, gy1));
    if (!(iw > 0.0f && ih > 0.0f)) return 0.0f;
    const float inter = K7_FMUL(iw, ih);
    float ua = K7_FSUB(K7_FADD(area_a, area_g), inter);
    ua = fmaxf(ua, 1e-8f);
    return K7_FDIV(inter, ua);
}

// ------------------------------------------------------------------ focal loss elements
#define K7_P_LO 1e-4f
#define K7_P_HI 0.9999f /* float32(1.0 - 1e-4), losses.py:62 */

// -log(1 - p) the way the reference gets it: q = fl(1 - p), then log(q).  For q >= 0.75
// it is evaluated as pp * S(pp) with pp = 1 - q (exact) and S(x) = -log(1-x)/x ~ a degree-5
// minimax polynomial on [0, 0.25] (max relative error 1.0e-7); otherwise libm logf.
K7_HD float k7_neg_log1m_poly(float pp) {
    float s = 0.3357100188732147f;
    s = fmaf(s, pp, 0.1467030793428421f);
    s = fmaf(s, pp, 0.25711026787757874f);
    s = fmaf(s, pp, 0.33293071389198303f);
    s = fmaf(s, pp, 0.5000072717666626f);
    s = fmaf(s, pp, 1.0f);
    return pp * s;
}
#define K7_POLY_MAX 0.25f

K7_HD float k7_neg_log1m(float q) {  // -log(q), q = fl(1-p)
    const float pp = 1.0f - q;
    return (pp <= K7_POLY_MAX) ? k7_neg_log1m_poly(pp) : -logf(q);
}

// Element with target 0 (negative anchor, or non-label class of a positive anchor):
//   loss = (1-alpha) * p^2 * -log(1-p)                                       losses.py:105-112
K7_HD float k7_focal_neg(float p, float nl, float one_m_alpha) { return (one_m_alpha * (p * p)) * nl; }
// d loss / d p = (1-alpha) * p * (2*nl + p/(1-p))
K7_HD float k7_focal_neg_grad(float p, float q, float nl, float one_m_alpha) {
#if defined(__CUDA_ARCH__)
    const float r = __fdividef(p, q);
#else
    const float r = p / q;
#endif
    return one_m_alpha * p * fmaf(2.0f, nl, r);
}
// Element with target 1 (label class of a positive anchor):
//   loss = alpha * (1-p)^2 * -log(p)
K7_HD float k7_focal_pos(float p, float alpha) {
    const float q = 1.0f - p;
    return (alpha * (q * q)) * (-logf(p));
}
K7_HD float k7_focal_pos_grad(float p, float alpha) {
    const float q = 1.0f - p;
    return -alpha * q * fmaf(2.0f, -logf(p), q / p);
}
K7_HD float k7_clamp_p(float x) { return fminf(fmaxf(x, K7_P_LO), K7_P_HI); }
// torch.sigmoid in float32 (ATen: 1 / (1 + exp(-x)) with expf and an IEEE division)
K7_HD float k7_sigmoid_precise(float z) { return K7_FDIV(1.0f, K7_FADD(1.0f, expf(-z))); }
// torch.clamp backward passes the gradient on the closed interval
K7_HD bool k7_clamp_pass(float x) { return x >= K7_P_LO && x <= K7_P_HI; }

// ------------------------------------------------------------------ regression targets + smooth L1
struct K7Reg4 {
    float v[4];
};

// Targets of a positive anchor (losses.py:125-148).  Anchor/GT in (x1,y1,x2,y2).
// out order (dx,dy,dw,dh) or (dy,dx,dh,dw) when yx_targets; divided by div[] when div != null.
K7_HD K7Reg4 k7_reg_targets(float ax1, float ay1, float ax2, float ay2, float gx1, float gy1, float gx2,
                            float gy2, const float* div, int yx_targets) {
    const float aw = K7_FSUB(ax2, ax1), ah = K7_FSUB(ay2, ay1);
    const float acx = K7_FADD(ax1, K7_FMUL(0.5f, aw)), acy = K7_FADD(ay1, K7_FMUL(0.5f, ah));
    float gw = K7_FSUB(gx2, gx1), gh = K7_FSUB(gy2, gy1);
    const float gcx = K7_FADD(gx1, K7_FMUL(0.5f, gw)), gcy = K7_FADD(gy1, K7_FMUL(0.5f, gh));
    gw = fmaxf(gw, 1.0f);
    gh = fmaxf(gh, 1.0f);
    const float dx = K7_FDIV(K7_FSUB(gcx, acx), aw);
    const float dy = K7_FDIV(K7_FSUB(gcy, acy), ah);
    const float dw = logf(K7_FDIV(gw, aw));
    const float dh = logf(K7_FDIV(gh, ah));
    K7Reg4 t;
    if (yx_targets) {
        t.v[0] = dy; t.v[1] = dx; t.v[2] = dh; t.v[3] = dw;
    } else {
        t.v[0] = dx; t.v[1] = dy; t.v[2] = dw; t.v[3] = dh;
    }
    if (div) {
#pragma unroll
        for (int i = 0; i < 4; ++i) t.v[i] = K7_FDIV(t.v[i], div[i]);
    }
    return t;
}

// smooth L1 of one coordinate (losses.py:155-160); *grad = d loss / d regression
K7_HD float k7_smooth_l1(float target, float reg, float beta, float quad, float shift, float* grad) {
    const float diff = K7_FSUB(target, reg);
    const float d = fabsf(diff);
    const bool small = d <= beta;
    if (grad) {
        const float sgn = (diff > 0.0f) ? -1.0f : ((diff < 0.0f) ? 1.0f : 0.0f);
        *grad = sgn * (small ? (2.0f * quad) * d : 1.0f);
    }
    return small ? K7_FMUL(quad, K7_FMUL(d, d)) : K7_FSUB(d, shift);
}

// ------------------------------------------------------------------ decode + clip (inference side)
struct K7Box {
    float x1, y1, x2, y2;
};

// BBoxTransform.forward, utils.py:26-53.  Anchor (x1,y1,x2,y2), deltas (dx,dy,dw,dh).
K7_HD K7Box k7_decode_retina(float ax1, float ay1, float ax2, float ay2, float r0, float r1, float r2,
                             float r3, const float* mean, const float* std) {
    const float w = K7_FSUB(ax2, ax1), h = K7_FSUB(ay2, ay1);
    const float cx = K7_FADD(ax1, K7_FMUL(0.5f, w)), cy = K7_FADD(ay1, K7_FMUL(0.5f, h));
    const float dx = K7_FADD(K7_FMUL(r0, std[0]), mean[0]);
    const float dy = K7_FADD(K7_FMUL(r1, std[1]), mean[1]);
    const float dw = K7_FADD(K7_FMUL(r2, std[2]), mean[2]);
    const float dh = K7_FADD(K7_FMUL(r3, std[3]), mean[3]);
    const float pcx = K7_FADD(cx, K7_FMUL(dx, w)), pcy = K7_FADD(cy, K7_FMUL(dy, h));
    const float pw = K7_FMUL(expf(dw), w), ph = K7_FMUL(expf(dh), h);
    K7Box b;
    b.x1 = K7_FSUB(pcx, K7_FMUL(0.5f, pw));
    b.y1 = K7_FSUB(pcy, K7_FMUL(0.5f, ph));
    b.x2 = K7_FADD(pcx, K7_FMUL(0.5f, pw));
    b.y2 = K7_FADD(pcy, K7_FMUL(0.5f, ph));
    return b;
}

// BBoxTransform.forward, utils_efficientdet.py:8-35.  Anchor given as (x1,y1,x2,y2) of the
// (y1,x1,y2,x2) tensor, deltas (dy,dx,dh,dw) -> box (x1,y1,x2,y2).
K7_HD K7Box k7_decode_effdet(float ax1, float ay1, float ax2, float ay2, float r0, float r1, float r2,
                             float r3) {
    const float yc = K7_FDIV(K7_FADD(ay1, ay2), 2.0f), xc = K7_FDIV(K7_FADD(ax1, ax2), 2.0f);
    const float ha = K7_FSUB(ay2, ay1), wa = K7_FSUB(ax2, ax1);
    const float w = K7_FMUL(expf(r3), wa), h = K7_FMUL(expf(r2), ha);
    const float y = K7_FADD(K7_FMUL(r0, ha), yc), x = K7_FADD(K7_FMUL(r1, wa), xc);
    K7Box b;
    b.x1 = K7_FSUB(x, K7_FDIV(w, 2.0f));
    b.y1 = K7_FSUB(y, K7_FDIV(h, 2.0f));
    b.x2 = K7_FADD(x, K7_FDIV(w, 2.0f));
    b.y2 = K7_FADD(y, K7_FDIV(h, 2.0f));
    return b;
}

// ClipBoxes.forward, utils.py:61-71: one-sided per coordinate.
K7_HD K7Box k7_clip_box(K7Box b, int height, int width) {
    b.x1 = fmaxf(b.x1, 0.0f);
    b.y1 = fmaxf(b.y1, 0.0f);
    b.x2 = fminf(b.x2, (float)(width - 1));
    b.y2 = fminf(b.y2, (float)(height - 1));
    return b;
}

// ------------------------------------------------------------------ NMS overlap predicate
// torchvision.ops.nms: suppress j when inter / (area_i + area_j - inter) > thr (NaN -> keep).
K7_HD bool k7_nms_suppress(const K7Box& a, float area_a, const K7Box& b, float area_b, float thr) {
    const float w = fmaxf(0.0f, K7_FSUB(fminf(a.x2, b.x2), fmaxf(a.x1, b.x1)));
    const float h = fmaxf(0.0f, K7_FSUB(fminf(a.y2, b.y2), fmaxf(a.y1, b.y1)));
    const float inter = K7_FMUL(w, h);
    const float ovr = K7_FDIV(inter, K7_FSUB(K7_FADD(area_a, area_b), inter));
    return ovr > thr;
}

// Sort key: scores descending, index ascending (stable descending sort of torchvision.ops.nms).
// Monotone map of IEEE float bits to unsigned, inverted.
K7_HD uint32_t k7_score_key_desc(float s) {
    s = K7_FADD(s, 0.0f);   // -0.0 -> +0.0: the two compare equal in torch's sort and must tie
#if defined(__CUDA_ARCH__)
    uint32_t u = __float_as_uint(s);
#else
    union { float f; uint32_t u; } cv; cv.f = s; uint32_t u = cv.u;
#endif
    u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    return ~u;
}
