// Focal + smooth-L1 loss (and its gradient) in one streaming pass (training side, stage 2).
//
//   k7_loss   FocalLoss.forward after assignment   architectures/losses.py:62, 94-168
//             (+ loss_efficientdet.py:61-160 through K7LossParams) and its autograd backward.
//
// One persistent kernel streams classifications [B*A, C] through per-warp TMA rings
// (k7_stream.cuh).  Per row the lane(s) owning it read the packed assignment word (staged next
// to the tile), evaluate every class as a "target 0" element with the polynomial path of
// k7_math.cuh, then patch the single "target 1" element of positive rows, so the inner loop is
// branch-free and identical for all rows.  With gradients, d(loss)/d(cls) overwrites the tile in
// shared memory and leaves through a bulk shared->global copy; d(loss)/d(reg) is one float4 per
// row.  Loss sums are reduced warp -> tile partial (one float4 per tile, written unconditionally,
// no atomics) and then per image in float64 in fixed order by the finalize kernel, so results
// are bit-reproducible run to run and independent of the grid size.
//
// Algorithmic bytes per image: forward A*(4C+16)+20M, forward+backward A*(8C+32)+20M.
#include "k7_stream.cuh"

#define K7_LOSS_MAX_THREADS 768

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

template <int VEC>
struct K7Vec;
template <>
struct K7Vec<1> {
    typedef float T;
};
template <>
struct K7Vec<2> {
    typedef float2 T;
};
template <>
struct K7Vec<4> {
    typedef float4 T;
};

// One vector of VEC classes of a non-ignored row: returns the loss contribution, optionally
// overwrites the values with gradients.
template <int VEC, bool GRAD>
__device__ __forceinline__ float k7_focal_vec(float* __restrict__ ptr, float one_m_alpha, float gscale) {
    typedef typename K7Vec<VEC>::T V;
    V raw = *reinterpret_cast<const V*>(ptr);
    float x[VEC], p[VEC], q[VEC], pp[VEC], nl[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) x[e] = reinterpret_cast<const float*>(&raw)[e];
    bool slow = false;
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
        p[e] = k7_clamp_p(x[e]);
        q[e] = 1.0f - p[e];
        pp[e] = 1.0f - q[e];
        nl[e] = k7_neg_log1m_poly(pp[e]);
        slow |= pp[e] > K7_POLY_MAX;
    }
    if (slow) {
#pragma unroll
        for (int e = 0; e < VEC; ++e)
            if (pp[e] > K7_POLY_MAX) nl[e] = -logf(q[e]);
    }
    float acc = 0.0f;
#pragma unroll
    for (int e = 0; e < VEC; ++e) acc = fmaf(one_m_alpha * (p[e] * p[e]), nl[e], acc);
    if (GRAD) {
        V out;
#pragma unroll
        for (int e = 0; e < VEC; ++e) {
            const float g = k7_focal_neg_grad(p[e], q[e], nl[e], one_m_alpha) * gscale;
            reinterpret_cast<float*>(&out)[e] = k7_clamp_pass(x[e]) ? g : 0.0f;
        }
        *reinterpret_cast<V*>(ptr) = out;
    }
    return acc;
}

template <int VEC, bool GRAD, bool FROM_SPEC>
__global__ void __launch_bounds__(K7_LOSS_MAX_THREADS, 1)
k7_loss_kernel(const __grid_constant__ K7AnchorSpec spec, const float* __restrict__ anchors, int box_order,
               const float* __restrict__ cls, const float* __restrict__ reg, const float* __restrict__ annotations,
               const int32_t* __restrict__ assign, const float2* __restrict__ scales,
               const __grid_constant__ K7LossDev P, const __grid_constant__ K7StreamGeom G,
               float4* __restrict__ partials, float* __restrict__ dcls, float4* __restrict__ dreg) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int NS = G.stages;
    // per-warp region: NS stages, then NS mbarriers (kept at the end of the CTA's shared memory)
    unsigned char* ring = smem + (size_t)warp * NS * G.stage_stride;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)G.warps * NS * G.stage_stride) + warp * NS;

    const long long gw = (long long)blockIdx.x * G.warps + warp;
    const long long nW = (long long)gridDim.x * G.warps;
    if (gw >= G.num_tiles) return;

    if (lane == 0) {
        for (int s = 0; s < NS; ++s) k7_mbar_init(bars + s, 1);
        k7_fence_mbar_init();
        k7_fence_proxy_async();
    }
    __syncwarp();

    // prefetch distance: with gradients a stage is load -> compute -> store, keep one stage for
    // the draining store; without, every stage but the current one can be in flight.
    const int PF = GRAD ? NS - 2 : NS - 1;
    if (lane == 0) {
        for (int k = 0; k < PF; ++k) {
            const long long t = gw + k * nW;
            if (t < G.num_tiles) k7_stream_issue(G, cls, assign, t, ring + (size_t)(k % NS) * G.stage_stride, bars + (k % NS));
        }
    }

    const int q = lane & (G.lpr - 1);
    const int rsub = lane >> G.lpr_shift;
    int clo, chi;
    k7_class_range(G, q, clo, chi);
    const float oma = P.one_m_alpha;

    int it = 0;
    for (long long tile = gw; tile < G.num_tiles; tile += nW, ++it) {
        const int s = it % NS;
        unsigned char* stage = ring + (size_t)s * G.stage_stride;
        // keep the ring full
        if (lane == 0) {
            const long long t = tile + (long long)PF * nW;
            if (t < G.num_tiles) {
                const int sn = (it + PF) % NS;
                if (GRAD) k7_bulk_wait_read<1>();   // the store that last used stage sn has been read out
                k7_stream_issue(G, cls, assign, t, ring + (size_t)sn * G.stage_stride, bars + sn);
            }
        }
        k7_mbar_wait(bars + s, (it / NS) & 1);

        const long long row0 = tile * (long long)G.T;
        long long rows_ll = G.total_rows - row0;
        const int rows = rows_ll > G.T ? G.T : (int)rows_ll;
        float* tilef = reinterpret_cast<float*>(stage);
        const int32_t* asg = reinterpret_cast<const int32_t*>(stage + G.tile_bytes);
        const uint32_t cls_bytes = (uint32_t)rows * G.C * 4u;
        const uint32_t rem = cls_bytes & 15u;
        if (rem) {   // only the last tile of a tensor with B*A*C % 4 != 0
            if (lane == 0) {
                const uint32_t base = (cls_bytes & ~15u) >> 2;
                for (uint32_t k = 0; k < (rem >> 2); ++k) tilef[base + k] = cls[row0 * G.C + base + k];
            }
            __syncwarp();
        }

        const int b0 = (int)(row0 / P.A);
        const long long split = (long long)(b0 + 1) * P.A;    // first row of image b0+1
        const float2 sc0 = __ldg(scales + b0);
        const float2 sc1 = __ldg(scales + min(b0 + 1, P.B - 1));
        float cls0 = 0.f, cls1 = 0.f, reg0 = 0.f, reg1 = 0.f;

        for (int i = 0; i < G.iters; ++i) {
            const int r = i * G.rows_per_iter + rsub;
            if (r >= rows) break;
            const long long R = row0 + r;
            const bool second = R >= split;
            const float2 sc = second ? sc1 : sc0;
            const int aw = asg[r];
            const int state = K7_ASSIGN_STATE(aw);
            float* rowp = tilef + (size_t)r * G.C;
            float rs = 0.0f;
            if (state != K7_STATE_IGNORE) {
                int label = -1;
                float xl = 0.0f;
                const int b = second ? b0 + 1 : b0;
                const float* arow = annotations + ((size_t)b * P.M + K7_ASSIGN_ROW(aw)) * 5;
                if (state == K7_STATE_POS) {
                    label = (int)__ldg(arow + 4);   // .long() truncation, losses.py:97
                    if (label >= clo && label < chi) xl = rowp[label]; else label = -1;
                }
                float a0 = 0.f, a1 = 0.f;
                int j = clo;
                for (; j + 2 * VEC <= chi; j += 2 * VEC) {
                    a0 += k7_focal_vec<VEC, GRAD>(rowp + j, oma, sc.x);
                    a1 += k7_focal_vec<VEC, GRAD>(rowp + j + VEC, oma, sc.x);
                }
                for (; j < chi; j += VEC) a0 += k7_focal_vec<VEC, GRAD>(rowp + j, oma, sc.x);
                rs = a0 + a1;
                if (label >= 0) {
                    // the label class has target 1: replace its "target 0" term
                    const float p = k7_clamp_p(xl);
                    const float qq = 1.0f - p;
                    const float nl = k7_neg_log1m(qq);
                    rs += k7_focal_pos(p, P.alpha) - k7_focal_neg(p, nl, oma);
                    if (GRAD) rowp[label] = k7_clamp_pass(xl) ? k7_focal_pos_grad(p, P.alpha) * sc.x : 0.0f;
                }
                // smooth-L1 on positives, one lane per row
                if (q == 0) {
                    float4 gr = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (state == K7_STATE_POS) {
                        const int a = (int)(R - (long long)b * P.A);
                        float ax1, ay1, ax2, ay2;
                        k7_anchor<FROM_SPEC>(spec, anchors, box_order, a, ax1, ay1, ax2, ay2);
                        const float g0 = __ldg(arow + 0), g1 = __ldg(arow + 1), g2 = __ldg(arow + 2), g3 = __ldg(arow + 3);
                        const K7Reg4 t = k7_reg_targets(ax1, ay1, ax2, ay2, g0, g1, g2, g3, P.use_div ? P.div : nullptr,
                                                        P.yx_targets);
                        const float4 rv = __ldg(reinterpret_cast<const float4*>(reg) + R);
                        float gx, gy, gz, gw_;
                        float l = k7_smooth_l1(t.v[0], rv.x, P.beta, P.quad, P.shift, &gx);
                        l += k7_smooth_l1(t.v[1], rv.y, P.beta, P.quad, P.shift, &gy);
                        l += k7_smooth_l1(t.v[2], rv.z, P.beta, P.quad, P.shift, &gz);
                        l += k7_smooth_l1(t.v[3], rv.w, P.beta, P.quad, P.shift, &gw_);
                        if (second) reg1 += l; else reg0 += l;
                        gr = make_float4(gx * sc.y, gy * sc.y, gz * sc.y, gw_ * sc.y);
                    }
                    if (GRAD) dreg[R] = gr;
                }
            } else if (GRAD) {
                typedef typename K7Vec<VEC>::T V;
                V z;
#pragma unroll
                for (int e = 0; e < VEC; ++e) reinterpret_cast<float*>(&z)[e] = 0.0f;
                for (int j = clo; j < chi; j += VEC) *reinterpret_cast<V*>(rowp + j) = z;
                if (q == 0) dreg[R] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            if (second) cls1 += rs; else cls0 += rs;
        }

        // tile partial: every tile writes its slot (no memset, no atomics)
        cls0 = k7_warp_sum(cls0);
        reg0 = k7_warp_sum(reg0);
        if (row0 + rows > split) {   // warp-uniform: the tile straddles two images
            cls1 = k7_warp_sum(cls1);
            reg1 = k7_warp_sum(reg1);
        }
        if (lane == 0) partials[tile] = make_float4(cls0, cls1, reg0, reg1);

        if (GRAD) {
            k7_fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                float* dst = dcls + row0 * G.C;
                const uint32_t bulk = cls_bytes & ~15u;
                if (bulk) k7_bulk_s2g(dst, stage, bulk);
                k7_bulk_commit();
                if (rem) {
                    const uint32_t base = bulk >> 2;
                    for (uint32_t k = 0; k < (rem >> 2); ++k) dst[base + k] = tilef[base + k];
                }
            }
        } else {
            __syncwarp();
        }
    }
    if (GRAD && lane == 0) k7_bulk_wait<0>();
}

// Per image: sum the tile partials in float64, fixed order; write (cls_loss_i, reg_loss_i).
// The last block to finish reduces the batch (threadfence + counter).
__global__ void __launch_bounds__(256)
k7_loss_finalize_kernel(const float4* __restrict__ partials, const int32_t* __restrict__ num_pos, int B, int A, int T,
                        float* __restrict__ per_image, float* __restrict__ sums, unsigned int* __restrict__ counter) {
    __shared__ double s_c[256], s_r[256];
    __shared__ bool s_last;
    const int b = blockIdx.x;
    const long long first = (long long)b * A, last = first + A - 1;
    const long long t0 = first / T, t1 = last / T;
    double c = 0.0, r = 0.0;
    for (long long t = t0 + threadIdx.x; t <= t1; t += blockDim.x) {
        const float4 p = partials[t];
        const bool second = (t * (long long)T) < first;   // tile started in image b-1
        c += second ? p.y : p.x;
        r += second ? p.w : p.z;
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
        __shared__ double s_n[256];
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
static size_t k7_loss_ws_layout(int B, long long num_tiles, size_t* off_scales, size_t* off_partials,
                                size_t* off_counter) {
    size_t o = 0;
    *off_counter = o;
    o += 256;
    *off_scales = o;
    o += k7_align_up(sizeof(float2) * (size_t)B, 256);
    *off_partials = o;
    o += k7_align_up(sizeof(float4) * (size_t)num_tiles, 256);
    return o;
}

extern "C" size_t k7_loss_workspace_bytes(int B, int A, int C) {
    if (B <= 0 || A <= 0 || C <= 0) return 0;
    // worst case for the tile count: the smallest tile the geometry can pick is 4 rows
    const long long rows = (long long)B * A;
    size_t a, b, c;
    return k7_loss_ws_layout(B, (rows + 3) / 4, &a, &b, &c);
}

template <int VEC, bool GRAD, bool FROM_SPEC>
static int k7_loss_launch(const K7AnchorSpec& spec, const float* anchors, int box_order, const float* cls,
                          const float* reg, const float* annotations, const int32_t* assign, const float2* scales,
                          const K7LossDev& P, const K7StreamGeom& G, float4* partials, float* dcls, float* dreg,
                          cudaStream_t st) {
    auto kern = k7_loss_kernel<VEC, GRAD, FROM_SPEC>;
    K7_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G.smem_bytes));
    kern<<<G.grid, G.warps * 32, G.smem_bytes, st>>>(spec, anchors, box_order, cls, reg, annotations, assign, scales, P, G,
                                                     partials, dcls, reinterpret_cast<float4*>(dreg));
    K7_LAUNCH_CHECK();
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
    K7_REQUIRE(B > 0 && A >= 32 && C > 0 && M > 0, "k7_loss: bad shape B=%d A=%d C=%d M=%d (A >= 32 required)", B, A, C, M);
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

    K7StreamGeom G;
    int rc = k7_stream_geometry((long long)B * A, A, C, 4, grad ? 4 : 3, K7_LOSS_MAX_THREADS / 32, &G);
    if (rc) return rc;

    size_t off_scales, off_partials, off_counter;
    const size_t need = k7_loss_ws_layout(B, G.num_tiles, &off_scales, &off_partials, &off_counter);
    if (workspace_bytes < need) {
        k7_set_error("k7_loss: workspace %zu < %zu bytes", workspace_bytes, need);
        return K7_EWORKSPACE;
    }
    unsigned char* ws = static_cast<unsigned char*>(workspace);
    K7_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 255) == 0, "k7_loss: workspace must be 256-byte aligned");
    float2* scales = reinterpret_cast<float2*>(ws + off_scales);
    float4* partials = reinterpret_cast<float4*>(ws + off_partials);
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
    k7_loss_scales_kernel<<<(B + 127) / 128, 128, 0, st>>>(num_pos, B, (float)params->mean_divisor, scales);
    K7_LAUNCH_CHECK();

    static const K7AnchorSpec kEmpty = {};
    const K7AnchorSpec& sp = spec ? *spec : kEmpty;
#define K7_LOSS_DISPATCH(V)                                                                                              \
    do {                                                                                                                 \
        if (grad) {                                                                                                      \
            rc = spec ? k7_loss_launch<V, true, true>(sp, anchors, box_order, cls, reg, annotations, assign, scales, P, G, \
                                                      partials, dcls, dreg, st)                                          \
                      : k7_loss_launch<V, true, false>(sp, anchors, box_order, cls, reg, annotations, assign, scales, P, \
                                                       G, partials, dcls, dreg, st);                                     \
        } else {                                                                                                         \
            rc = spec ? k7_loss_launch<V, false, true>(sp, anchors, box_order, cls, reg, annotations, assign, scales, P, \
                                                       G, partials, dcls, dreg, st)                                      \
                      : k7_loss_launch<V, false, false>(sp, anchors, box_order, cls, reg, annotations, assign, scales,   \
                                                        P, G, partials, dcls, dreg, st);                                 \
        }                                                                                                                \
    } while (0)
    if (G.vec == 4)
        K7_LOSS_DISPATCH(4);
    else if (G.vec == 2)
        K7_LOSS_DISPATCH(2);
    else
        K7_LOSS_DISPATCH(1);
#undef K7_LOSS_DISPATCH
    if (rc) return rc;

    k7_loss_finalize_kernel<<<B, 256, 0, st>>>(partials, num_pos, B, A, G.T, per_image, sums, counter);
    K7_LAUNCH_CHECK();
    return K7_OK;
}
