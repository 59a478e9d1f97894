// Library-level pieces of the C ABI: version, error string, argument validation, the streaming
// geometry chooser.
#include <stdlib.h>
#include <string.h>

#include "k7_stream.cuh"

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

static int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    if (!v || !*v) return dflt;
    return atoi(v);
}

int k7_stream_geometry(long long total_rows, int A, int C, int side_bytes_per_row, int stages_default, int max_warps,
                       K7StreamGeom* g) {
    memset(g, 0, sizeof(*g));
    g->C = C;
    g->vec = (C % 4 == 0) ? 4 : ((C % 2 == 0) ? 2 : 1);
    int lpr = 1;
    while (lpr < 32 && (C + lpr - 1) / lpr > 32) lpr <<= 1;
    // a lane's class range must hold at least one vector
    while (lpr > 1 && C / lpr < g->vec) lpr >>= 1;
    g->lpr = lpr;
    g->lpr_shift = 0;
    while ((1 << g->lpr_shift) < lpr) ++g->lpr_shift;
    g->rows_per_iter = 32 / lpr;
    const int row_bytes = C * 4;
    const int tile_target = env_int("K7_TILE_BYTES", 4096);
    // iterations per tile: as close to the byte target as possible, T % 4 == 0, T <= A
    int unit = g->rows_per_iter;          // rows added per iteration
    int iters_step = 1;
    while ((unit * iters_step) % 4 != 0) ++iters_step;
    int iters = (int)((double)tile_target / ((double)unit * row_bytes) + 0.5);
    iters = iters / iters_step * iters_step;
    if (iters < iters_step) iters = iters_step;
    while (iters > iters_step && (long long)unit * iters > A) iters -= iters_step;
    K7_REQUIRE((long long)unit * iters <= A, "streaming geometry: A=%d is smaller than the minimum tile of %d rows", A,
               unit * iters);
    g->iters = iters;
    g->T = unit * iters;
    g->tile_bytes = (uint32_t)g->T * row_bytes;
    g->side_bytes = (uint32_t)g->T * side_bytes_per_row;
    g->stage_stride = (uint32_t)k7_align_up((size_t)g->tile_bytes + g->side_bytes, 128);
    g->stages = env_int("K7_STAGES", stages_default);
    if (g->stages < 3) g->stages = 3;
    if (g->stages > 8) g->stages = 8;
    const size_t budget = 200 * 1024;
    const size_t per_warp = (size_t)g->stages * g->stage_stride + (size_t)g->stages * 8;
    K7_REQUIRE(per_warp <= budget, "streaming geometry: C=%d needs %zu bytes of shared memory per warp", C, per_warp);
    int warps = (int)(budget / per_warp);
    const int warps_cap = env_int("K7_WARPS", max_warps);
    if (warps > warps_cap) warps = warps_cap;
    if (warps > max_warps) warps = max_warps;
    if (warps < 1) warps = 1;
    g->warps = warps;
    g->smem_bytes = (size_t)warps * per_warp;
    g->total_rows = total_rows;
    g->num_tiles = (total_rows + g->T - 1) / g->T;
    long long ctas_needed = (g->num_tiles + warps - 1) / warps;
    int grid = k7_sm_count();
    if (ctas_needed < grid) grid = (int)ctas_needed;
    if (grid < 1) grid = 1;
    g->grid = grid;
    return K7_OK;
}
