// Warp-autonomous streaming of the [B*A, C] classification tensor through shared memory.
//
// Both hot kernels (loss, detect) read every class probability exactly once and do a few ALU
// operations per element, so the whole job is "move A*C*4 bytes per image at HBM speed".
// The tensor is viewed as a flat list of R = B*A rows of C floats.  A tile is T consecutive rows
// (T*C*4 bytes, contiguous in memory, 16-byte aligned because T % 4 == 0).  Each WARP owns a
// private ring of NS shared-memory stages with one mbarrier per stage; lane 0 issues 1-D bulk
// async copies (TMA, SASS UBLKCP) global->shared for tiles ahead of the one being computed, the
// warp waits on the stage's mbarrier, computes with lane <-> (row, class-range) mapping, and (loss
// with gradients) writes results in place and ships the stage back with a bulk shared->global
// copy.  No __syncthreads anywhere: warps never wait for each other, tiles are assigned
// round-robin over all warps of the persistent grid (one CTA per SM).
//
// Lane mapping: `lpr` lanes share a row (lpr = power of two, chosen so a lane's class range is
// <= 32 elements), a warp covers 32/lpr rows per iteration and `iters` iterations per tile.
// Lane q of a row handles classes [lo_q, hi_q) with boundaries rounded to the vector width so
// every shared-memory access is a naturally aligned LDS.32/64/128.
#pragma once

#include "k7_common.cuh"

struct K7StreamGeom {
    int C;               // classes (floats per row)
    int vec;             // 1, 2 or 4 floats per shared-memory access
    int lpr;             // lanes per row (power of two)
    int lpr_shift;       // log2(lpr)
    int rows_per_iter;   // 32 / lpr
    int iters;           // iterations per tile
    int T;               // rows per tile = rows_per_iter * iters, multiple of 4
    int stages;          // NS
    int warps;           // warps per CTA
    int grid;            // CTAs
    uint32_t tile_bytes;     // T*C*4
    uint32_t side_bytes;     // per-row side data staged with the tile (T*4: assignment words) or 0
    uint32_t stage_stride;   // bytes between stages, multiple of 128
    size_t smem_bytes;       // dynamic shared memory per CTA
    long long total_rows;    // B*A
    long long num_tiles;
};

// Host: choose the geometry.  `side` = bytes of per-row side data staged with each tile.
// tile_target / warps_max can be overridden by the environment for tuning sweeps:
//   K7_TILE_BYTES, K7_WARPS, K7_STAGES.
int k7_stream_geometry(long long total_rows, int A, int C, int side_bytes_per_row, int stages_default,
                       int max_warps, K7StreamGeom* g);

__device__ __forceinline__ void k7_class_range(const K7StreamGeom& g, int q, int& lo, int& hi) {
    lo = ((q * g.C) >> g.lpr_shift) / g.vec * g.vec;
    hi = (q == g.lpr - 1) ? g.C : ((((q + 1) * g.C) >> g.lpr_shift) / g.vec * g.vec);
}

// Issue the loads of tile `tile` into stage memory `dst` (lane 0 only).
//   cls_bytes = rows*C*4 of the tile; the bulk copy moves the 16-byte multiple, the caller's
//   lane 0 moves the (<= 12 byte) remainder after the wait (only the last tile of a tensor whose
//   element count is not a multiple of 4 has one).
__device__ __forceinline__ void k7_stream_issue(const K7StreamGeom& g, const float* __restrict__ cls,
                                                const int32_t* __restrict__ side, long long tile,
                                                unsigned char* dst, uint64_t* bar) {
    const long long row0 = tile * (long long)g.T;
    long long rows = g.total_rows - row0;
    if (rows > g.T) rows = g.T;
    const uint32_t cls_bytes = (uint32_t)(rows * g.C * 4);
    const uint32_t bulk = cls_bytes & ~15u;
    uint32_t side_bulk = 0;
    if (side) side_bulk = ((uint32_t)(rows * 4) + 15u) & ~15u;   // side buffer is padded by the caller
    if (bulk + side_bulk) {
        k7_mbar_expect_tx(bar, bulk + side_bulk);
        if (bulk) k7_bulk_g2s(dst, cls + row0 * g.C, bulk, bar);
        if (side_bulk) k7_bulk_g2s(dst + g.tile_bytes, side + row0, side_bulk, bar);
    } else {
        k7_mbar_arrive(bar);
    }
}
