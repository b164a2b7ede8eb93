// primitives.cu -- device-wide exclusive scan and stable LSD radix sort of (key,val) pairs.
//
// These are the deterministic building blocks of scale-down victim order (README.md:214:
// "the first k VMs in list order") and of placement admission order: both
// need "rank among earlier rows of the same bucket", i.e. a STABLE partition by bucket id.
// No atomics decide any output position, so results do not depend on scheduling order.
#include <mutex>

#include "prims.cuh"

namespace k8sgpu {

// ------------------------------------------------------------------ exclusive scan (u32)

__global__ void __launch_bounds__(kScanThreads) scan_reduce_kernel(const uint32_t *__restrict__ in, uint32_t n,
                                                                   uint32_t *__restrict__ block_sums) {
    uint32_t total = block_scan_tile(in, nullptr, blockIdx.x * kScanTile, n, 0u, false);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// (in and out may be the same buffer -- the allocation-free entry point documents in-place use --
// so neither is __restrict__: every thread loads its eight elements before it stores them.)
__global__ void __launch_bounds__(kScanThreads) scan_down_kernel(const uint32_t *in, uint32_t *out, uint32_t n,
                                                                 const uint32_t *__restrict__ block_offs) {
    block_scan_tile(in, out, blockIdx.x * kScanTile, n, block_offs[blockIdx.x], true);
}

// Single block of 1024 threads: scans any n by looping over 8192-element tiles; writes total.
constexpr int kScanSingleThreads = 1024;
__global__ void __launch_bounds__(kScanSingleThreads) scan_single_kernel(const uint32_t *in, uint32_t *out, uint32_t n,
                                                                         uint32_t *__restrict__ total) {
    uint32_t carry = 0;
    for (uint32_t base = 0; base < n; base += kScanSingleThreads * 8)
        carry = block_scan_tile<kScanSingleThreads>(in, out, base, n, carry, true);
    if (threadIdx.x == 0 && total) *total = carry;
}

int exclusive_scan_u32(k8sgpu_ctx *ctx, cudaStream_t s, const uint32_t *d_in, uint32_t *d_out, uint32_t n,
                       DevBuf &tmp, uint32_t *d_total) {
    if (n <= 8 * kScanTile) {  // small: one block walks it
        K8S_LAUNCH(ctx, scan_single_kernel, 1, kScanSingleThreads, 0, s, d_in, d_out, n, d_total);
        return K8SGPU_OK;
    }
    uint32_t nb = ceil_div(n, kScanTile);
    // tmp holds the block sums of this level; deeper levels use the tail of the same buffer
    K8S_CUDA(ctx, tmp.reserve(((size_t)nb * 2 + 64) * sizeof(uint32_t)));
    uint32_t *sums = tmp.as<uint32_t>();
    K8S_LAUNCH(ctx, scan_reduce_kernel, nb, kScanThreads, 0, s, d_in, n, sums);
    // nb <= 2^31 / 2048 = 2^20 block sums: one block scans them in place
    K8S_LAUNCH(ctx, scan_single_kernel, 1, kScanSingleThreads, 0, s, sums, sums, nb, d_total);
    K8S_LAUNCH(ctx, scan_down_kernel, nb, kScanThreads, 0, s, d_in, d_out, n, sums);
    return K8SGPU_OK;
}

// ------------------------------------------------------------------ stable LSD radix sort

// Per-block digit histogram, written digit-major: hist[d * nblocks + block].
__global__ void __launch_bounds__(kSortThreads) sort_hist_kernel(const uint32_t *__restrict__ keys, uint32_t n,
                                                                 int shift, uint32_t mask,
                                                                 uint32_t *__restrict__ hist, uint32_t nblocks) {
    sort_hist_tile(keys, n, shift, mask, hist, nblocks, blockIdx.x);
}

// Stable scatter of one tile (prims.cuh).
__global__ void __launch_bounds__(kSortThreads) sort_scatter_kernel(
    const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, uint32_t *__restrict__ out_keys,
    uint32_t *__restrict__ out_vals, uint32_t n, int shift, uint32_t mask, const uint32_t *__restrict__ hist_scan,
    uint32_t nblocks) {
    sort_scatter_tile(keys, vals, out_keys, out_vals, n, shift, mask, hist_scan, nblocks, blockIdx.x);
}

int sort_pairs(k8sgpu_ctx *ctx, cudaStream_t s, uint32_t *keys, uint32_t *vals, uint32_t n, int bits,
               SortScratch &scr) {
    if (n <= 1 || bits <= 0) return K8SGPU_OK;
    const uint32_t nb = ceil_div(n, kSortTile);
    K8S_CUDA(ctx, scr.keys_alt.reserve((size_t)n * 4));
    K8S_CUDA(ctx, scr.vals_alt.reserve((size_t)n * 4));
    K8S_CUDA(ctx, scr.hist.reserve(((size_t)nb * 256 * 2 + 256) * 4));
    uint32_t *hist = scr.hist.as<uint32_t>();
    uint32_t *hist_scan = hist + (size_t)nb * 256;

    uint32_t *src_k = keys, *src_v = vals, *dst_k = scr.keys_alt.as<uint32_t>(), *dst_v = scr.vals_alt.as<uint32_t>();
    int passes = 0;
    for (int shift = 0; shift < bits; shift += 8, ++passes) {
        int w = bits - shift < 8 ? bits - shift : 8;
        uint32_t mask = (1u << w) - 1u;
        K8S_LAUNCH(ctx, sort_hist_kernel, nb, kSortThreads, 0, s, src_k, n, shift, mask, hist, nb);
        int rc = exclusive_scan_u32(ctx, s, hist, hist_scan, nb * 256, scr.scan_tmp, nullptr);
        if (rc) return rc;
        K8S_LAUNCH(ctx, sort_scatter_kernel, nb, kSortThreads, 0, s, src_k, src_v, dst_k, dst_v, n, shift, mask,
                   hist_scan, nb);
        uint32_t *t;
        t = src_k; src_k = dst_k; dst_k = t;
        t = src_v; src_v = dst_v; dst_v = t;
    }
    if (passes & 1) {
        K8S_CUDA(ctx, cudaMemcpyAsync(keys, src_k, (size_t)n * 4, cudaMemcpyDeviceToDevice, s));
        K8S_CUDA(ctx, cudaMemcpyAsync(vals, src_v, (size_t)n * 4, cudaMemcpyDeviceToDevice, s));
    }
    return K8SGPU_OK;
}

}  // namespace k8sgpu

// ------------------------------------------------------------------ exclusive scan, no allocation
//
// Who this is for: a caller that must not allocate device memory while it works.  Two kinds
// exist around this library.  One is a scheduler or controller process that shares its GPU with
// the jobs it places and was given a fixed slice of device memory at start-up: a later
// cudaMalloc may simply fail there.  The other is a caller in the middle of queued GPU work it
// cannot afford to have stalled: cudaMalloc and cudaFree synchronise the whole device.
//
// The scan needs one word of scratch per 2048-element block (the block sums).  Here that scratch
// is a fixed array in device global memory, part of the library image: every GPU that loads
// libk8sgpu.so holds these 32 KiB until the process exits.  The size caps the input at
// K8SGPU_SCAN_NOALLOC_MAX_N elements.
namespace k8sgpu {

constexpr uint32_t kNoAllocScanMaxN = K8SGPU_SCAN_NOALLOC_MAX_N;
constexpr uint32_t kNoAllocScanBlocks = kNoAllocScanMaxN / kScanTile;  // 8192 block sums = 32 KiB
static __device__ __align__(256) uint32_t g_noalloc_block_sums[kNoAllocScanBlocks];

constexpr int kNoAllocMaxDevices = 64;
// Device address of g_noalloc_block_sums on each GPU.  On the host the array's name is only a
// handle for the runtime, not something a kernel can dereference, so the address is looked up
// once per device (with that device current) and cached here.
static uint32_t *g_noalloc_ptr[kNoAllocMaxDevices] = {};
static std::mutex g_noalloc_mu;

static int noalloc_scratch(k8sgpu_ctx *ctx, uint32_t **out) {
    if (ctx->device < 0 || ctx->device >= kNoAllocMaxDevices)
        return fail(ctx, K8SGPU_E_ARG, "scan_noalloc: device %d not supported (max %d)", ctx->device,
                    kNoAllocMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[ctx->device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_block_sums);  // K8S_ENTER made ctx->device current
        if (e != cudaSuccess || !p)
            return fail(ctx, K8SGPU_E_CUDA, "scan_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        g_noalloc_ptr[ctx->device] = static_cast<uint32_t *>(p);
    }
    *out = g_noalloc_ptr[ctx->device];
    return K8SGPU_OK;
}

}  // namespace k8sgpu

extern "C" int k8sgpu_exclusive_scan_noalloc(k8sgpu_ctx *ctx, const uint32_t *d_in, uint32_t n, uint32_t *d_out,
                                             uint32_t *d_total, void *stream) {
    using namespace k8sgpu;
    K8S_ENTER(ctx);
    if (n > kNoAllocScanMaxN)
        return fail(ctx, K8SGPU_E_CAPACITY,
                    "scan_noalloc: n=%u exceeds the built-in scratch (at most %u elements); use the allocating "
                    "entry points for larger inputs",
                    n, kNoAllocScanMaxN);
    if (n && (!d_in || !d_out)) return fail(ctx, K8SGPU_E_ARG, "scan_noalloc: NULL device buffer");
    cudaStream_t s = ::k8sgpu::pick_stream(ctx, stream);
    uint32_t *sums = nullptr;
    int rc = noalloc_scratch(ctx, &sums);
    if (rc) return rc;
    if (n <= kScanTile) {  // one block, no scratch touched
        K8S_LAUNCH(ctx, scan_single_kernel, 1, kScanSingleThreads, 0, s, d_in, d_out, n, d_total);
        return K8SGPU_OK;
    }
    // Nothing in the scratch has to start at zero: the first kernel overwrites every block sum
    // that the later two read, so whatever the previous call left behind is never seen.
    const uint32_t nb = ceil_div(n, kScanTile);
    K8S_LAUNCH(ctx, scan_reduce_kernel, nb, kScanThreads, 0, s, d_in, n, sums);
    K8S_LAUNCH(ctx, scan_single_kernel, 1, kScanSingleThreads, 0, s, sums, sums, nb, d_total);
    K8S_LAUNCH(ctx, scan_down_kernel, nb, kScanThreads, 0, s, d_in, d_out, n, sums);
    return K8SGPU_OK;
}
