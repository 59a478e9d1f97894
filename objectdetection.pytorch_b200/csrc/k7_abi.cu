// Library-level pieces of the C ABI: version, error string, argument validation.
#include <stdlib.h>
#include <string.h>

#include "k7_common.cuh"

static thread_local char g_err[512] = "";

void k7_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* k7_last_error(void) { return g_err; }
extern "C" int k7_version(void) { return K7_ABI_VERSION; }

int k7_sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

int k7_check_spec(const K7AnchorSpec* s, int A) {
    K7_REQUIRE(s != nullptr, "anchor spec is null");
    K7_REQUIRE(s->num_levels >= 1 && s->num_levels <= K7_MAX_LEVELS, "anchor spec: num_levels=%d out of range", s->num_levels);
    K7_REQUIRE(s->cell_anchors >= 1 && s->cell_anchors <= K7_MAX_CELL_ANCHORS, "anchor spec: cell_anchors=%d out of range",
               s->cell_anchors);
    K7_REQUIRE(s->box_order == K7_BOX_XYXY || s->box_order == K7_BOX_YXYX, "anchor spec: bad box_order");
    long long off = 0;
    for (int l = 0; l < s->num_levels; ++l) {
        K7_REQUIRE(s->fh[l] >= 0 && s->fw[l] >= 0, "anchor spec: negative grid at level %d", l);
        K7_REQUIRE(s->offset[l] == off, "anchor spec: offset[%d]=%d, expected %lld", l, s->offset[l], off);
        off += (long long)s->fh[l] * s->fw[l] * s->cell_anchors;
    }
    K7_REQUIRE(off == s->total && s->offset[s->num_levels] == s->total, "anchor spec: total=%d, levels sum to %lld",
               s->total, off);
    K7_REQUIRE(s->total == A, "anchor spec describes %d anchors but the tensors have A=%d", s->total, A);
    return K7_OK;
}

