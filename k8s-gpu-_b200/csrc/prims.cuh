// prims.cuh -- per-tile device functions of the scan and the stable radix sort.  They are written
// against a *virtual* tile index so that the same code serves the one-tile-per-CTA kernels in
// primitives.cu and the persistent cooperative admission kernel in place.cu.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace k8sgpu {

namespace cg = cooperative_groups;

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t x, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t y = __shfl_up_sync(0xFFFFFFFFu, x, d);
        if (lane >= d) x += y;
    }
    return x;
}

// Scans one tile of kScanTile elements with 256 threads: each thread owns 8 consecutive
// elements (two 128-bit loads when the tile is whole and 16-byte aligned), scans them in
// registers, and one warp-shuffle + shared-memory step combines the 256 thread totals.
// carry_in is added to every output.  Returns carry_in + the tile total (all threads).
// THREADS is the CTA size (256, or 1024 for the one-CTA kernel); a tile is THREADS * 8 elements.
template <int THREADS = 256>
__device__ __forceinline__ uint32_t block_scan_tile(const uint32_t *in, uint32_t *out,  // may alias (in place)
                                                    uint32_t base, uint32_t n, uint32_t carry_in, bool write) {
    constexpr int NW = THREADS / 32;
    __shared__ uint32_t warp_tot[NW];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t i0 = base + (uint32_t)tid * 8u;
    const bool vec = base + (uint32_t)THREADS * 8u <= n &&
                     (((uintptr_t)in | (write ? (uintptr_t)out : (uintptr_t)0)) & 15u) == 0;
    uint32_t x[8];
    if (vec) {
        const uint4 a = *reinterpret_cast<const uint4 *>(in + i0);
        const uint4 b = *reinterpret_cast<const uint4 *>(in + i0 + 4);
        x[0] = a.x; x[1] = a.y; x[2] = a.z; x[3] = a.w;
        x[4] = b.x; x[5] = b.y; x[6] = b.z; x[7] = b.w;
    } else {
#pragma unroll
        for (int k = 0; k < 8; ++k) x[k] = (i0 + k < n) ? in[i0 + k] : 0u;
    }
#pragma unroll
    for (int k = 1; k < 8; ++k) x[k] += x[k - 1];
    const uint32_t inc = warp_incl_scan(x[7], lane);
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t woff = 0, all = 0;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        const uint32_t t = warp_tot[w];
        if (w < warp) woff += t;
        all += t;
    }
    if (write) {
        const uint32_t e = carry_in + woff + inc - x[7];  // exclusive prefix of this thread's first element
        if (vec) {
            *reinterpret_cast<uint4 *>(out + i0) = make_uint4(e, e + x[0], e + x[1], e + x[2]);
            *reinterpret_cast<uint4 *>(out + i0 + 4) = make_uint4(e + x[3], e + x[4], e + x[5], e + x[6]);
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (i0 + k < n) out[i0 + k] = e + (k ? x[k - 1] : 0u);
        }
    }
    __syncthreads();  // warp_tot is reused by the next tile
    return carry_in + all;
}

// Per-block digit histogram, written digit-major: hist[d * nblocks + block].
__device__ __forceinline__ void sort_hist_tile(const uint32_t *__restrict__ keys, uint32_t n, int shift, uint32_t mask,
                                               uint32_t *__restrict__ hist, uint32_t nblocks, uint32_t vt) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = vt * kSortTile;
#pragma unroll
    for (uint32_t r = 0; r < kSortTile / kSortThreads; ++r) {
        uint32_t i = base + r * kSortThreads + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & mask], 1u);  // counts only: order-free
    }
    __syncthreads();
    hist[threadIdx.x * nblocks + vt] = h[threadIdx.x];
    __syncthreads();  // h is reused by the caller's next tile
}

// Stable scatter.  Warp w owns the contiguous 256-element chunk w of the tile and walks it in
// 8 sub-rows of 32, so (warp, sub-row, lane) order == input order.  Ranks inside a sub-row come
// from __match_any_sync; per-warp digit counters carry the rank across sub-rows.  The tile is
// then laid out digit-major in shared memory and written out by consecutive threads, so each
// (block, digit) run leaves as one contiguous store instead of 4-byte scattered writes.
//
// hist_scan is the exclusive scan of the digit-major histogram -- or, with local_scan, the raw
// histogram itself: thread d then sums row d (all tiles of digit d) and the row's prefix up to this
// tile, and a block scan over the 256 row totals gives the same number.  Every CTA re-reads the
// whole histogram (256 * nblocks words, L2 resident), which is cheaper than a separate grid-wide
// scan phase plus its barrier while nblocks is small (kLocalScanMaxTiles).
constexpr uint32_t kLocalScanMaxTiles = 64;
__device__ __forceinline__ void sort_scatter_tile(
    const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, uint32_t *__restrict__ out_keys,
    uint32_t *__restrict__ out_vals, uint32_t n, int shift, uint32_t mask, const uint32_t *__restrict__ hist_scan,
    uint32_t nblocks, uint32_t vt, bool local_scan = false) {
    __shared__ uint32_t cnt[8][256];
    __shared__ uint32_t gdelta[256];  // global base of digit d minus its start inside the tile
    __shared__ uint32_t wtot[8], wtot2[8];
    __shared__ uint32_t sk[kSortTile], sv[kSortTile];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int w = 0; w < 8; ++w) cnt[w][tid] = 0;
    __syncthreads();

    const uint32_t tile_base = vt * kSortTile;
    const uint32_t chunk = tile_base + warp * 256;
    uint32_t k[8], v[8], rank[8];
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        uint32_t i = chunk + s * 32 + lane;
        bool in = i < n;
        k[s] = in ? keys[i] : 0u;
        v[s] = in ? vals[i] : 0u;
    }
    uint32_t gbase = 0, row_tot = 0;
    if (local_scan) {  // issued before the ranking so the loads overlap it
        const uint32_t *row = hist_scan + (size_t)tid * nblocks;
#pragma unroll 8
        for (uint32_t b = 0; b < nblocks; ++b) {
            const uint32_t h = row[b];
            gbase += b < vt ? h : 0u;
            row_tot += h;
        }
    } else {
        gbase = hist_scan[(size_t)tid * nblocks + vt];
    }
    const uint32_t lt = (1u << lane) - 1u;
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        uint32_t i = chunk + s * 32 + lane;
        bool in = i < n;
        uint32_t live = __ballot_sync(0xFFFFFFFFu, in);
        rank[s] = 0;
        if (in) {
            uint32_t d = (k[s] >> shift) & mask;
            uint32_t peers = __match_any_sync(live, d);  // (8 ballots, one per digit bit, measured the same)
            uint32_t before = cnt[warp][d];
            rank[s] = before + __popc(peers & lt);
            __syncwarp(live);
            if ((peers & lt) == 0) cnt[warp][d] = before + __popc(peers);  // lowest peer updates
        }
        __syncwarp();
    }
    __syncthreads();
    {  // thread d: per-warp counts of digit d -> offsets inside the digit's run; run starts by block scan
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            uint32_t t = cnt[w][tid];
            cnt[w][tid] = run;
            run += t;
        }
        uint32_t inc = warp_incl_scan(run, lane);  // exclusive start of digit `tid` inside the tile
        uint32_t inc2 = local_scan ? warp_incl_scan(row_tot, lane) : 0u;  // ... and in the whole array
        if (lane == 31) {
            wtot[warp] = inc;
            wtot2[warp] = inc2;
        }
        __syncthreads();
        uint32_t woff = 0, woff2 = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w)
            if (w < warp) {
                woff += wtot[w];
                woff2 += wtot2[w];
            }
        const uint32_t dstart = woff + inc - run;
#pragma unroll
        for (int w = 0; w < 8; ++w) cnt[w][tid] += dstart;
        if (local_scan) gbase += woff2 + inc2 - row_tot;
        gdelta[tid] = gbase - dstart;
    }
    __syncthreads();
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        uint32_t i = chunk + s * 32 + lane;
        if (i < n) {
            uint32_t d = (k[s] >> shift) & mask;
            uint32_t p = cnt[warp][d] + rank[s];
            sk[p] = k[s];
            sv[p] = v[s];
        }
    }
    __syncthreads();
    const uint32_t count = min(kSortTile, n - tile_base);
#pragma unroll
    for (int s = 0; s < 8; ++s) {
        uint32_t p = s * kSortThreads + tid;
        if (p < count) {
            uint32_t kk = sk[p];
            uint32_t dst = gdelta[(kk >> shift) & mask] + p;
            out_keys[dst] = kk;
            out_vals[dst] = sv[p];
        }
    }
    __syncthreads();  // shared staging is reused by the caller's next tile
}

// Exclusive scan of n u32 by the whole grid; every CTA must call it.  Ends with a grid barrier.
__device__ __forceinline__ void coop_exclusive_scan(cg::grid_group &grid, const uint32_t *in, uint32_t *out, uint32_t n,
                                                    uint32_t *sums, uint32_t *total) {
    const uint32_t nt = (n + kScanTile - 1) / kScanTile;
    if (nt <= 16) {  // small: CTA 0 walks it
        if (blockIdx.x == 0) {
            uint32_t carry = 0;
            for (uint32_t t = 0; t < nt; ++t) carry = block_scan_tile(in, out, t * kScanTile, n, carry, true);
            if (threadIdx.x == 0 && total) *total = carry;
        }
        grid.sync();
        return;
    }
    for (uint32_t t = blockIdx.x; t < nt; t += gridDim.x) {
        uint32_t tot = block_scan_tile(in, nullptr, t * kScanTile, n, 0u, false);
        if (threadIdx.x == 0) sums[t] = tot;
    }
    grid.sync();
    if (blockIdx.x == 0) {  // nt <= 2^31 / 2048: one CTA scans the tile sums in place
        uint32_t carry = 0;
        for (uint32_t t = 0; t * kScanTile < nt; ++t) carry = block_scan_tile(sums, sums, t * kScanTile, nt, carry, true);
        if (threadIdx.x == 0 && total) *total = carry;
    }
    grid.sync();
    for (uint32_t t = blockIdx.x; t < nt; t += gridDim.x) block_scan_tile(in, out, t * kScanTile, n, sums[t], true);
    grid.sync();
}

}  // namespace k8sgpu
