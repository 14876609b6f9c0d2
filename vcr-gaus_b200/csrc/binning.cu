// binning.cu — Gaussian -> tile instance lists: prefix sums, depth sort of the Gaussians, instance emission in depth order,
// stable LSD radix sort by tile id, per-tile ranges, tile schedule.  All hand-written (no CUB/thrust).
//
// Replaces, in the reference (DGR/cuda_rasterizer/rasterizer_impl.cu):
//   cub::DeviceScan::InclusiveSum            :283      -> preprocess block sums + scan_block_sums_kernel (-> R), then
//                                                         sorted_block_sums_kernel + scan for the depth-ordered emission
//   duplicateWithKeys                        :70-111   -> duplicate_sorted_kernel (block-cooperative, coalesced; emits in
//                                                         depth order).  duplicate_kernel = the reference's index order,
//                                                         used by the parity view only
//   cub::DeviceRadixSort::SortPairs          :309-314  -> radix_pass_kernel: 4 passes over the P (depth, index) pairs +
//                                                         ceil(tile_bits / 8) passes over the R (tile, index) pairs; both
//                                                         as (u32 key, u32 value), 8 bytes per pair (the reference moves
//                                                         12-byte (u64, u32) pairs through 6 passes)
//   cudaMemset + identifyTileRanges          :316-323  -> tile_ranges_kernel (+ tile_order_kernel, the block schedule)
//
// The sorted point list and the tile ranges are bit-identical to the reference's by construction: a stable sort by tile of
// an emission that is already in (depth, index) order yields, inside every tile, the order (depth, index) — which is what
// CUB's stable sort of (tile << 32 | depth) keys over the reference's index-ordered emission yields.  The reference's
// intermediate lists (index-ordered duplicated keys, point_offsets, 64-bit sorted keys) are produced on demand by
// vcr_launch_reference_lists for the parity tests.
#include "common.cuh"
#include <atomic>

// A look-back that never finds its predecessor (a block that died, a corrupted status array) must not hang the GPU: the
// spin is bounded, and a block that gives up raises this per-device flag and carries on with a wrong offset.  The flag
// is collected by scan_block_sums_kernel into the 16 bytes the NEXT forward reads back together with R, where it becomes
// an error (vcr_rasterize_forward: "radix sort look-back timed out"); a debug forward checks it before returning.
// While the flag is up the kernels that consume the (now invalid) lists — duplicate_sorted_kernel, tile_ranges_kernel and the
// compositing kernels of render_forward.cu / render_backward.cu, which get the flag's address as an argument — return at
// once instead of indexing with garbage; the outputs of that call are then simply unwritten.
static __device__ unsigned int g_radix_timeout = 0u;

#include <mutex>
const unsigned int* vcr_radix_abort_flag() {        // device address of the flag on the current device (cached per device)
    static const unsigned int* cache[64] = {nullptr};
    static std::mutex mtx;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lock(mtx);
    if (!cache[dev]) {
        void* p = nullptr;
        if (cudaGetSymbolAddress(&p, g_radix_timeout) != cudaSuccess) return nullptr;
        cache[dev] = static_cast<const unsigned int*>(p);
    }
    return cache[dev];
}

// ------------------------------------------------------------------------------------------------
// Prefix sum over the per-block totals written by preprocess_kernel (one block, 1024 threads).
// block_sums[b] becomes the exclusive prefix; num_rendered[0] the grand total.  With take_error the kernel also moves
// the radix timeout flag into num_rendered[3] (the words the host reads back together with R).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) scan_block_sums_kernel(uint32_t* __restrict__ block_sums, int nb,
                                                               uint32_t* __restrict__ num_rendered, int take_error) {
    __shared__ uint32_t warp_tot[32];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    // every thread owns `ipt` consecutive entries: one block-wide scan of the per-thread totals instead of one per 1024 entries
    const int ipt = (nb + 1023) / 1024;
    const int first = t * ipt, last = min(first + ipt, nb);
    uint32_t v = 0;
    for (int i = first; i < last; i++) v += block_sums[i];
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) warp_tot[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t x = warp_tot[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += n;
        }
        warp_tot[lane] = x;  // inclusive over warps
    }
    __syncthreads();
    uint32_t run = (w ? warp_tot[w - 1] : 0u) + inc - v;     // exclusive prefix of this thread's first entry
    for (int i = first; i < last; i++) {
        const uint32_t c = block_sums[i];
        block_sums[i] = run;
        run += c;
    }
    if (t == 1023) {
        num_rendered[0] = warp_tot[31];
        if (take_error) num_rendered[3] = atomicExch(&g_radix_timeout, 0u);
    }
}

int vcr_launch_scan_block_sums(int P, GeomState& g, cudaStream_t s) {
    const int nb = (P + 255) / 256;
    VcrProfScope prof(VCR_K_SCAN, s);
    scan_block_sums_kernel<<<1, 1024, 0, s>>>(g.block_sums, nb, g.num_rendered, 1);
    VCR_LAUNCH_CHECK("scan_block_sums_kernel");
    return 0;
}

// ------------------------------------------------------------------------------------------------
// point_offsets + key duplication + radix digit histograms, one block per 256 Gaussians.
// The block's instances occupy the contiguous output range [base, base + total); the 256 threads
// walk that range together (coalesced 8-byte key / 4-byte value stores) and find the owning
// Gaussian of each slot by binary search in the block's inclusive scan.
// ------------------------------------------------------------------------------------------------
#define VCR_MAX_PASSES 8
__global__ void __launch_bounds__(256) duplicate_kernel(int P, int gx, int gy, const int* __restrict__ radii,
                                                        const GeomState g, uint64_t* __restrict__ keys_out,
                                                        uint32_t* __restrict__ vals_out, uint32_t* __restrict__ ghist,
                                                        int npass, int key_bits) {
    __shared__ uint32_t incl[256];
    __shared__ uint32_t s_depth[256];
    __shared__ int s_x0[256], s_y0[256], s_w[256];
    __shared__ uint32_t warp_tot[8];
    __shared__ uint32_t hist[VCR_MAX_PASSES][256];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int idx = blockIdx.x * 256 + t;
    for (int k = t; k < npass * 256; k += 256) (&hist[0][0])[k] = 0u;

    uint32_t tiles = 0, dbits = 0;
    int x0 = 0, y0 = 0, rw = 0;
    if (idx < P) {
        tiles = g.tiles_touched[idx];
        if (tiles > 0) {
            int2 rmin, rmax;
            vcr_get_rect(g.means2D[idx], radii[idx], gx, gy, rmin, rmax);
            x0 = rmin.x; y0 = rmin.y; rw = rmax.x - rmin.x;
            dbits = __float_as_uint(g.depths[idx]);
        }
    }
    uint32_t inc = tiles;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) warp_tot[w] = inc;
    __syncthreads();
    uint32_t wex = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) wex += (k < w) ? warp_tot[k] : 0u;
    inc += wex;
    incl[t] = inc;
    s_depth[t] = dbits; s_x0[t] = x0; s_y0[t] = y0; s_w[t] = rw;
    const uint32_t base = g.block_sums[blockIdx.x];
    if (idx < P) g.point_offsets[idx] = base + inc;
    // depth digits: one histogram update per Gaussian (weight = its instance count)
    if (tiles > 0) {
        const int ndepth = npass < 4 ? npass : 4;
        for (int p = 0; p < ndepth; p++) atomicAdd(&hist[p][(dbits >> (8 * p)) & 255u], tiles);
    }
    __syncthreads();
    const uint32_t total = incl[255];
    for (uint32_t k = t; k < total; k += 256) {
        // smallest gi with incl[gi] > k
        int lo = 0, hi = 255;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (incl[mid] > k) hi = mid; else lo = mid + 1;
        }
        const uint32_t start = lo ? incl[lo - 1] : 0u;
        const uint32_t local = k - start;
        const int wdt = s_w[lo];
        const int y = s_y0[lo] + (int)(local / (uint32_t)wdt);
        const int x = s_x0[lo] + (int)(local % (uint32_t)wdt);
        const uint32_t tile = (uint32_t)(y * gx + x);
        keys_out[base + k] = ((uint64_t)tile << 32) | s_depth[lo];
        vals_out[base + k] = (uint32_t)(blockIdx.x * 256 + lo);
        for (int p = 4; p < npass; p++) {
            const int nb = min(8, key_bits - 8 * p);
            atomicAdd(&hist[p][(tile >> (8 * (p - 4))) & ((1u << nb) - 1u)], 1u);
        }
    }
    __syncthreads();
    for (int p = 0; p < npass; p++) {
        const uint32_t c = hist[p][t];
        if (c) atomicAdd(&ghist[p * 256 + t], c);
    }
}

// ------------------------------------------------------------------------------------------------
// Production emission: instances in DEPTH order.  Sorting the P Gaussians by depth first (vcr_launch_gaussian_sort) and
// emitting their instances in that order leaves only the tile id to sort by (2 passes for up to 65536 tiles instead of
// 6 passes over 44-45 key bits); the result is the same list the reference's (tile | depth) sort produces.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sorted_block_sums_kernel(int P, const uint32_t* __restrict__ sorted_ids,
                                                                const uint32_t* __restrict__ tiles_touched,
                                                                uint32_t* __restrict__ sums) {
    __shared__ uint32_t warp_tot[8];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (g_radix_timeout) return;                    // failed depth sort: sorted_ids is garbage
    const uint32_t v = i < P ? tiles_touched[sorted_ids[i]] : 0u;
    const uint32_t wsum = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31) == 0) warp_tot[threadIdx.x >> 5] = wsum;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) t += warp_tot[w];
        sums[blockIdx.x] = t;
    }
}

__global__ void __launch_bounds__(256) duplicate_sorted_kernel(int P, int gx, int gy, const int* __restrict__ radii,
                                                               const GeomState g, const uint32_t* __restrict__ sorted_ids,
                                                               const uint32_t* __restrict__ sorted_sums,
                                                               uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                                                               uint32_t* __restrict__ ghist, int npass, int tile_bits) {
    __shared__ uint32_t incl[256];
    __shared__ uint32_t s_gid[256];
    __shared__ int s_x0[256], s_y0[256], s_w[256];
    __shared__ float s_iw[256];                     // 1 / rectangle width
    __shared__ uint32_t warp_tot[8];
    __shared__ uint32_t hist[4][256];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int i = blockIdx.x * 256 + t;
    if (g_radix_timeout) return;                    // the depth sort before this launch failed: sorted_ids is garbage
    for (int p = 0; p < npass; p++) hist[p][t] = 0u;

    uint32_t tiles = 0, gid = 0;
    int x0 = 0, y0 = 0, rw = 0;
    if (i < P) {
        gid = sorted_ids[i];
        tiles = g.tiles_touched[gid];
        if (tiles > 0) {
            int2 rmin, rmax;
            vcr_get_rect(g.means2D[gid], radii[gid], gx, gy, rmin, rmax);
            x0 = rmin.x; y0 = rmin.y; rw = rmax.x - rmin.x;
        }
    }
    uint32_t inc = tiles;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) warp_tot[w] = inc;
    __syncthreads();
    uint32_t wex = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) wex += (k < w) ? warp_tot[k] : 0u;
    inc += wex;
    incl[t] = inc;
    s_gid[t] = gid; s_x0[t] = x0; s_y0[t] = y0; s_w[t] = rw; s_iw[t] = rw > 0 ? 1.0f / (float)rw : 0.f;
    const uint32_t base = sorted_sums[blockIdx.x];
    __syncthreads();
    const uint32_t total = incl[255];
    for (uint32_t k = t; k < total; k += 256) {
        int lo = 0, hi = 255;                           // smallest entry with incl > k
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (incl[mid] > k) hi = mid; else lo = mid + 1;
        }
        const uint32_t local = k - (lo ? incl[lo - 1] : 0u);
        const int wdt = s_w[lo];
        // row = local / width.  (local + 0.5) * (1 / width) in FP32 has the same integer part as the exact quotient while
        // local < 2^22: the quotient's fraction lies in [0.5/w, 1 - 0.5/w] and the two roundings move it by < local * 2^-23 / w.
        // Rectangles of 2^22 tiles and more (a 32k x 32k image covered by one Gaussian) take the integer division.
        uint32_t row;
        if (local < (1u << 22)) row = (uint32_t)(((float)local + 0.5f) * s_iw[lo]);
        else row = local / (uint32_t)wdt;
        const int y = s_y0[lo] + (int)row;
        const int x = s_x0[lo] + (int)(local - row * (uint32_t)wdt);
        const uint32_t tile = (uint32_t)(y * gx + x);
        keys_out[base + k] = tile;
        vals_out[base + k] = s_gid[lo];
        for (int p = 0; p < npass; p++) {
            int sh, nb;
            vcr_radix_pass_bits(tile_bits, p, sh, nb);             // the digit widths vcr_launch_tile_sort uses
            atomicAdd(&hist[p][(tile >> sh) & ((1u << nb) - 1u)], 1u);
        }
    }
    __syncthreads();
    for (int p = 0; p < npass; p++) {
        const uint32_t c = hist[p][t];
        if (c) atomicAdd(&ghist[p * 256 + t], c);
    }
}

int vcr_launch_duplicate_sorted(const FwdParams& p, GeomState& g, BinningState& b, int R, int tile_bits, cudaStream_t s) {
    (void)R;
    const int npass = (tile_bits + 7) / 8;          // <= 4: tile ids are 32-bit
    const int nb = (p.P + 255) / 256;
    { VcrProfScope m(VCR_K_MEMSET, s); VCR_CUDA_TRY(cudaMemsetAsync(b.sort_hist, 0, VCR_MAX_PASSES * 256 * sizeof(uint32_t), s)); }
    {
        VcrProfScope prof(VCR_K_SCAN, s, 2);
        sorted_block_sums_kernel<<<nb, 256, 0, s>>>(p.P, g.gs_vals[0], g.tiles_touched, g.sorted_sums);
        VCR_LAUNCH_CHECK("sorted_block_sums_kernel");
        scan_block_sums_kernel<<<1, 1024, 0, s>>>(g.sorted_sums, nb, g.num_rendered + 2, 0);
        VCR_LAUNCH_CHECK("scan_block_sums_kernel");
    }
    VcrProfScope prof(VCR_K_DUPLICATE, s);
    // the sort ends in buffer A (tile_keys, point_list): an even number of passes may start there, an odd number starts in B
    uint32_t* ek = (npass & 1) ? b.alt_keys : b.tile_keys;
    uint32_t* ev = (npass & 1) ? b.alt_vals : b.point_list;
    duplicate_sorted_kernel<<<nb, 256, 0, s>>>(p.P, p.gx, p.gy, p.radii, g, g.gs_vals[0], g.sorted_sums, ek, ev, b.sort_hist, npass, tile_bits);
    VCR_LAUNCH_CHECK("duplicate_sorted_kernel");
    return 0;
}

// ------------------------------------------------------------------------------------------------
// Parity view (tests only): materialise the lists in the REFERENCE's formats from the production state —
//   point_list_keys_unsorted / point_list_unsorted : duplicateWithKeys' output, Gaussian-index order, (tile << 32 | depth)
//   point_offsets                                  : the inclusive scan of tiles_touched in index order
//   point_list_keys                                : the sorted keys as (tile << 32 | depth) (production keeps bare u32 tiles)
// The three arrays are separate from everything the production path reads, so the call can be repeated.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) reference_keys_kernel(uint32_t R, uint64_t* __restrict__ keys, const uint32_t* __restrict__ tiles,
                                                             const uint32_t* __restrict__ point_list, const float* __restrict__ depths) {
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i >= R) return;
    keys[i] = ((uint64_t)tiles[i] << 32) | (uint64_t)__float_as_uint(depths[point_list[i]]);
}

int vcr_launch_reference_lists(int P, int R, int gx, int gy, const int* radii, GeomState& g, BinningState& b, cudaStream_t s) {
    if (R <= 0) return 0;
    duplicate_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, gx, gy, radii, g, b.point_list_keys_unsorted, b.point_list_unsorted,
                                                     b.sort_hist, 0, 0);
    VCR_LAUNCH_CHECK("duplicate_kernel");
    reference_keys_kernel<<<((uint32_t)R + 255u) / 256u, 256, 0, s>>>((uint32_t)R, b.point_list_keys, b.tile_keys, b.point_list, g.depths);
    VCR_LAUNCH_CHECK("reference_keys_kernel");
    return 0;
}

// Exclusive scan of each pass's 256-bin histogram (one block per pass).
__global__ void __launch_bounds__(256) radix_scan_hist_kernel(uint32_t* __restrict__ ghist) {
    __shared__ uint32_t warp_tot[8];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    uint32_t* h = ghist + blockIdx.x * 256;
    const uint32_t v = h[t];
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) warp_tot[w] = inc;
    __syncthreads();
    uint32_t wex = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) wex += (k < w) ? warp_tot[k] : 0u;
    h[t] = wex + inc - v;
}

// ------------------------------------------------------------------------------------------------
// One stable radix pass (<= 8 bits) over R (u32 key, u32 value) pairs.
//   - tiles of 4096 pairs per 256-thread block, taken in ticket order (forward progress for look-back)
//   - in-block stable ranks: per-warp running digit counters + __match_any_sync peer masks
//   - cross-block digit offsets: decoupled look-back on (flag | count) words, one thread per digit
//   - pairs are first scattered into shared memory in block-sorted order so the global stores are
//     runs of consecutive addresses per digit
// ------------------------------------------------------------------------------------------------
#define RADIX_THREADS VCR_RADIX_THREADS
#define RADIX_WARPS (RADIX_THREADS / 32)
#define RADIX_ITEMS VCR_RADIX_ITEMS
#define RADIX_TILE (RADIX_THREADS * RADIX_ITEMS)
#define FLAG_LOCAL 0x40000000u
#define FLAG_GLOBAL 0x80000000u
#define FLAG_MASK 0xC0000000u
#define RADIX_SMEM_BYTES (RADIX_TILE * 4 + RADIX_TILE * 4 + RADIX_WARPS * 256 * 4 + 256 * 4 + 256 * 4 + 128)

static uint32_t g_radix_spin_limit = 1u << 26;      // vcr_debug_radix_fault() lowers it and silences block 0
static int g_radix_fault = 0;

template <int NB>      // ballots per item in the ranking: >= the digit width of the pass
__global__ void __launch_bounds__(RADIX_THREADS, VCR_RADIX_MINBLOCKS) radix_pass_kernel(const uint32_t* __restrict__ keys_in,
                                                                      const uint32_t* __restrict__ vals_in,
                                                                      uint32_t* __restrict__ keys_out,
                                                                      uint32_t* __restrict__ vals_out, uint32_t R, int shift,
                                                                      int nbits, const uint32_t* __restrict__ digit_base,
                                                                      volatile uint32_t* status, uint32_t* ticket,
                                                                      uint32_t spin_limit, int fault) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t* s_keys = reinterpret_cast<uint32_t*>(smem_raw);
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(smem_raw + RADIX_TILE * 4);
    uint32_t* s_whist = s_vals + RADIX_TILE;             // [RADIX_WARPS][256] per-warp running digit counters
    uint32_t* s_dstart = s_whist + RADIX_WARPS * 256;    // [256] block-local exclusive digit start
    uint32_t* s_gbase = s_dstart + 256;                  // [256] global base minus local start
    uint32_t* s_misc = s_gbase + 256;                    // [32]
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const uint32_t mask = (1u << nbits) - 1u;

    if (t == 0) s_misc[0] = atomicAdd(ticket, 1u);
#pragma unroll
    for (int k = 0; k < 8; k++) s_whist[w * 256 + lane + 32 * k] = 0u;
    __syncthreads();
    const uint32_t bid = s_misc[0];
    const uint32_t base = bid * RADIX_TILE;
    const uint32_t n = min((uint32_t)RADIX_TILE, R - base);

    uint32_t key[RADIX_ITEMS];
    uint32_t val[RADIX_ITEMS];
    uint32_t rank[RADIX_ITEMS];
#pragma unroll
    for (int i = 0; i < RADIX_ITEMS; i++) {
        const uint32_t li = w * (32 * RADIX_ITEMS) + i * 32 + lane;
        if (li < n) { key[i] = keys_in[base + li]; val[i] = vals_in[base + li]; }
        else { key[i] = ~0u; val[i] = 0u; }
    }
    uint32_t* my_hist = s_whist + w * 256;
#pragma unroll
    for (int i = 0; i < RADIX_ITEMS; i++) {
        const uint32_t li = w * (32 * RADIX_ITEMS) + i * 32 + lane;
        const bool valid = li < n;
        const uint32_t d = (key[i] >> shift) & mask;
        const uint32_t vmask = __ballot_sync(0xffffffffu, valid);
#ifndef VCR_RADIX_MATCH_ANY
        // peer mask from one ballot per digit bit: constant cost, whereas match.any costs one round per distinct digit in
        // the warp — and both the depth-ordered emission (adjacent tiles of one Gaussian) and the index-ordered depth keys
        // put ~30 distinct digits into a warp (89 us against 62 for the low tile byte)
        uint32_t peers = vmask;
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const uint32_t bit = (d >> b) & 1u;
            const uint32_t m = __ballot_sync(0xffffffffu, bit != 0u);
            peers &= bit ? m : ~m;
        }
#else
        const uint32_t peers = __match_any_sync(0xffffffffu, d) & vmask;
#endif
        uint32_t prev = 0;
        if (valid) prev = my_hist[d];
        rank[i] = prev + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
        if (valid && lane == (__ffs(peers) - 1)) my_hist[d] = prev + __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // threads 0..255 own one digit each
    uint32_t cnt = 0, inc = 0;
    volatile uint32_t* my_status = status + (size_t)bid * 256;
    if (t < 256) {
#pragma unroll
        for (int k = 0; k < RADIX_WARPS; k++) {
            const uint32_t c = s_whist[k * 256 + t];
            s_whist[k * 256 + t] = cnt;  // exclusive over warps
            cnt += c;
        }
        // Publish the block-local count as early as possible.  Flag and count share one 32-bit word, so no fence is
        // needed: a reader either sees the old word (keeps polling) or the complete new one.
        if (!(fault && bid == 0)) my_status[t] = (bid == 0 ? FLAG_GLOBAL : FLAG_LOCAL) | cnt;   // (fault: test hook, block 0 stays silent)
        // block-local exclusive scan over digits (8 warps)
        inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += x;
        }
        if (lane == 31) s_misc[1 + w] = inc;
    }
    __syncthreads();
    if (t < 256) {
        uint32_t wex = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) wex += (k < w) ? s_misc[1 + k] : 0u;
        const uint32_t dstart = wex + inc - cnt;
        s_dstart[t] = dstart;
        // decoupled look-back
        uint32_t excl = 0;
        if (bid > 0) {
            // Walk back over the predecessors' words, RB_LOOK of them in flight at a time: when all blocks of a pass are
            // resident at once (the P-sized depth sort: a few hundred blocks) block i has to look ~i/2 blocks back, and
            // one dependent L2 round trip per block made that walk the longest thing in the kernel.
#ifndef VCR_RADIX_LOOK
#define VCR_RADIX_LOOK 4
#endif
            constexpr int RB_LOOK = VCR_RADIX_LOOK;
            int j = (int)bid - 1;
            uint32_t spins = 0;
            bool found = false;
            while (!found) {
                uint32_t v[RB_LOOK];
#pragma unroll
                for (int k = 0; k < RB_LOOK; k++) v[k] = (j - k >= 0) ? status[(size_t)(j - k) * 256 + t] : 0u;
                int used = 0;
#pragma unroll
                for (int k = 0; k < RB_LOOK; k++) {
                    if (found || used != k) continue;            // stop at the first word that is not published yet
                    const uint32_t f = v[k] & FLAG_MASK;
                    if (f == 0u) continue;
                    excl += v[k] & ~FLAG_MASK;
                    used = k + 1;
                    if (f == FLAG_GLOBAL) found = true;
                }
                j -= used;
                if (!found && used == 0 && ++spins > spin_limit) { atomicExch(&g_radix_timeout, 1u); break; }
            }
            my_status[t] = FLAG_GLOBAL | (excl + cnt);
        }
        s_gbase[t] = digit_base[t] + excl - dstart;
    }
    __syncthreads();
    // scatter into shared memory in block-sorted order
#pragma unroll
    for (int i = 0; i < RADIX_ITEMS; i++) {
        const uint32_t li = w * (32 * RADIX_ITEMS) + i * 32 + lane;
        if (li < n) {
            const uint32_t d = (key[i] >> shift) & mask;
            const uint32_t slot = s_dstart[d] + my_hist[d] + rank[i];
            s_keys[slot] = key[i];
            s_vals[slot] = val[i];
        }
    }
    __syncthreads();
    for (uint32_t sl = t; sl < n; sl += RADIX_THREADS) {
        const uint32_t k = s_keys[sl];
        const uint32_t d = (k >> shift) & mask;
        const uint32_t pos = s_gbase[d] + sl;
        keys_out[pos] = k;
        vals_out[pos] = s_vals[sl];
    }
}

// Stable LSD sort of n (u32 key, u32 value) pairs over key bits [0, key_bits), 8 bits per pass.  `hist` holds the digit
// histogram of every pass ([npass][256], turned into exclusive offsets here); the sorted pairs end in (kA, vA), with
// (kB, vB) as the other ping-pong buffer.  (kA, vA) may alias the input when the number of passes is even.
template <int NB>
static void radix_pass_launch(uint32_t nblocks, cudaStream_t s, const uint32_t* kin, const uint32_t* vin, uint32_t* kout, uint32_t* vout, uint32_t n,
                              int shift, int nbits, const uint32_t* digit_base, uint32_t* status, uint32_t* ticket) {
    radix_pass_kernel<NB><<<nblocks, RADIX_THREADS, RADIX_SMEM_BYTES, s>>>(kin, vin, kout, vout, n, shift, nbits, digit_base, status, ticket,
                                                                           g_radix_spin_limit, g_radix_fault);
}
// `balanced`: ceil(bits/8) passes of (almost) equal width (vcr_radix_pass_bits) instead of 8-bit digits and a remainder
static int radix_sort_run(const uint32_t* kin, const uint32_t* vin, uint32_t* kA, uint32_t* vA, uint32_t* kB, uint32_t* vB,
                          uint32_t n, int key_bits, uint32_t* hist, uint32_t* status, cudaStream_t s, bool balanced = false) {
    // the opt-in shared-memory size is a per-device function attribute (setting it twice is harmless, so a benign race
    // between two host threads on the atomic flag only costs a repeated call)
    static std::atomic<bool> attr_set_dev[64];
    int dev_ = 0;
    VCR_CUDA_TRY(cudaGetDevice(&dev_));
    if (dev_ >= 0 && dev_ < 64 && !attr_set_dev[dev_].load(std::memory_order_acquire)) {
        VCR_CUDA_TRY(cudaFuncSetAttribute(radix_pass_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, RADIX_SMEM_BYTES));
        VCR_CUDA_TRY(cudaFuncSetAttribute(radix_pass_kernel<7>, cudaFuncAttributeMaxDynamicSharedMemorySize, RADIX_SMEM_BYTES));
        VCR_CUDA_TRY(cudaFuncSetAttribute(radix_pass_kernel<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, RADIX_SMEM_BYTES));
        VCR_CUDA_TRY(cudaFuncSetAttribute(radix_pass_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, RADIX_SMEM_BYTES));
        VCR_CUDA_TRY(cudaFuncSetAttribute(radix_pass_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, RADIX_SMEM_BYTES));
        attr_set_dev[dev_].store(true, std::memory_order_release);
    }
    if (key_bits < 1 || key_bits > 32) { vcr_set_error("radix sort: 1..32 key bits"); return 1; }
    if (n >= (1u << 30)) { vcr_set_error("radix sort: at most 2^30 - 1 pairs (the look-back words carry flag | count in 32 bits)"); return 1; }
    const int npass = (key_bits + 7) / 8;
    const uint32_t nblocks = (n + RADIX_TILE - 1) / RADIX_TILE;
    { VcrProfScope prof(VCR_K_RADIX_HIST, s); radix_scan_hist_kernel<<<npass, 256, 0, s>>>(hist); }
    VCR_LAUNCH_CHECK("radix_scan_hist_kernel");
    // status words for all passes + one ticket per pass + error flag, zeroed in one memset
    const size_t words = (size_t)npass * nblocks * 256 + 16;
    { VcrProfScope m(VCR_K_MEMSET, s); VCR_CUDA_TRY(cudaMemsetAsync(status, 0, words * sizeof(uint32_t), s)); }
    uint32_t* tickets = status + (size_t)npass * nblocks * 256;
    for (int pass = 0; pass < npass; pass++) {
        const bool to_final = ((npass - 1 - pass) % 2) == 0;
        uint32_t* kout = to_final ? kA : kB;
        uint32_t* vout = to_final ? vA : vB;
        int shift = 8 * pass, nbits = key_bits - 8 * pass < 8 ? key_bits - 8 * pass : 8;
        if (balanced) vcr_radix_pass_bits(key_bits, pass, shift, nbits);
        VcrProfScope prof(VCR_K_RADIX_PASS, s);
        uint32_t* st_ = status + (size_t)pass * nblocks * 256;
        if (nbits <= 4) radix_pass_launch<4>(nblocks, s, kin, vin, kout, vout, n, shift, nbits, hist + pass * 256, st_, tickets + pass);
        else if (nbits == 5) radix_pass_launch<5>(nblocks, s, kin, vin, kout, vout, n, shift, nbits, hist + pass * 256, st_, tickets + pass);
        else if (nbits == 6) radix_pass_launch<6>(nblocks, s, kin, vin, kout, vout, n, shift, nbits, hist + pass * 256, st_, tickets + pass);
        else if (nbits == 7) radix_pass_launch<7>(nblocks, s, kin, vin, kout, vout, n, shift, nbits, hist + pass * 256, st_, tickets + pass);
        else radix_pass_launch<8>(nblocks, s, kin, vin, kout, vout, n, shift, nbits, hist + pass * 256, st_, tickets + pass);
        VCR_LAUNCH_CHECK("radix_pass_kernel");
        kin = kout; vin = vout;
    }
    return 0;
}

// Generic entry (kNN's Morton sort, an even number of passes): sorts (tile_keys, point_list) in place, B as ping-pong.
int vcr_launch_radix_sort(BinningState& b, int R, int key_bits, cudaStream_t s) {
    if (((key_bits + 7) / 8) & 1) { vcr_set_error("vcr_launch_radix_sort: in-place sort needs an even number of 8-bit passes"); return 1; }
    return radix_sort_run(b.tile_keys, b.point_list, b.tile_keys, b.point_list, b.alt_keys, b.alt_vals, (uint32_t)R, key_bits, b.sort_hist,
                          b.sort_status, s);
}

// Move the look-back timeout flag of this device into *flag_out_device (and clear it), ordered on `s`.
__global__ void radix_take_error_kernel(uint32_t* out) { *out = atomicExch(&g_radix_timeout, 0u); }
int vcr_radix_take_error(uint32_t* flag_out_device, cudaStream_t s) {
    radix_take_error_kernel<<<1, 1, 0, s>>>(flag_out_device);
    VCR_LAUNCH_CHECK("radix_take_error_kernel");
    return 0;
}
// Test hook (tests/test_rasterizer_gpu.py): with on != 0 block 0 of every radix pass publishes nothing and the spin limit
// drops to 2000 polls, so every multi-block sort times out (and produces garbage) — the failure the flag exists to report.
extern "C" int vcr_debug_radix_fault(int on) {
    g_radix_fault = on ? 1 : 0;
    g_radix_spin_limit = on ? 2000u : (1u << 26);
    return 0;
}

// Sort ALL Gaussians by depth bits (stable, i.e. ties in index order): 4 passes over P pairs written by preprocess_kernel.
// Result in g.gs_keys[0] / g.gs_vals[0]; Gaussians that are not rendered (key 0xFFFFFFFF) end up behind the others.
int vcr_launch_gaussian_sort(int P, GeomState& g, cudaStream_t s) {
    return radix_sort_run(g.gs_keys[0], g.gs_vals[0], g.gs_keys[0], g.gs_vals[0], g.gs_keys[1], g.gs_vals[1], (uint32_t)P, 32, g.gs_hist,
                          g.gs_status, s);
}

// Stable sort of the R instances (emitted in depth order by duplicate_sorted_kernel) by tile id alone: ceil(tile_bits/8)
// passes.  Within a tile the depth order of the emission survives, which is exactly the order the reference gets from
// sorting (tile << 32 | depth) keys.  Input where vcr_launch_duplicate_sorted left it (buffer A for an even number of
// passes, B for an odd one), result in buffer A = (tile_keys, point_list).
int vcr_launch_tile_sort(BinningState& b, int R, int tile_bits, cudaStream_t s) {
    const bool odd = ((tile_bits + 7) / 8) & 1;
    return radix_sort_run(odd ? b.alt_keys : b.tile_keys, odd ? b.alt_vals : b.point_list, b.tile_keys, b.point_list, b.alt_keys, b.alt_vals,
                          (uint32_t)R, tile_bits, b.sort_hist, b.sort_status, s, true);
}

// ------------------------------------------------------------------------------------------------
// Tile ranges (identifyTileRanges, rasterizer_impl.cu:116-138).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) tile_ranges_kernel(uint32_t L, const uint32_t* __restrict__ keys, uint2* __restrict__ ranges) {
    // four consecutive keys per thread (one 16-byte load) and the key before them
    const uint32_t idx = 4u * (blockIdx.x * 256u + threadIdx.x);
    if (idx >= L || g_radix_timeout) return;           // (a failed sort leaves garbage tile ids: do not index with them)
    uint32_t k[4];                                     // the production keys are bare tile ids
    if (idx + 4u <= L) {
        const uint4 q = *reinterpret_cast<const uint4*>(keys + idx);
        k[0] = q.x; k[1] = q.y; k[2] = q.z; k[3] = q.w;
    } else {
#pragma unroll
        for (int e = 0; e < 4; e++) k[e] = idx + e < L ? keys[idx + e] : 0u;
    }
    uint32_t prev = idx ? keys[idx - 1] : 0u;
#pragma unroll
    for (int e = 0; e < 4; e++) {
        const uint32_t i = idx + e;
        if (i >= L) break;
        const uint32_t cur = k[e];
        if (i == 0) ranges[cur].x = 0;
        else if (cur != prev) { ranges[prev].y = i; ranges[cur].x = i; }
        if (i == L - 1) ranges[cur].y = L;
        prev = cur;
    }
}

// Block schedule for the compositing kernels: tile ids ordered by list length, longest first (coarse: by
// floor(log2(length)), arbitrary inside a bucket).  Per-tile list lengths are extremely skewed (median tens, maximum
// tens of thousands of entries), and thread blocks are dispatched in index order, so without this the longest tiles
// can start last and run alone at the end of the kernel.  One block; counting sort over 33 buckets.
__global__ void __launch_bounds__(1024) tile_order_kernel(const uint2* __restrict__ ranges, int tiles, uint32_t* __restrict__ order) {
    __shared__ uint32_t cnt[33];
    __shared__ uint32_t start[33];
    const int t = threadIdx.x;
    if (t < 33) cnt[t] = 0u;
    __syncthreads();
    for (int i = t; i < tiles; i += 1024) {
        const uint2 r = ranges[i];
        const uint32_t n = r.y - r.x;
        atomicAdd(&cnt[n ? 32 - __clz(n) : 0], 1u);     // bucket 0: empty tiles, bucket k: 2^(k-1) <= n < 2^k
    }
    __syncthreads();
    if (t == 0) {
        uint32_t acc = 0;
        for (int b = 32; b >= 0; b--) { start[b] = acc; acc += cnt[b]; }   // longest first
    }
    __syncthreads();
    for (int i = t; i < tiles; i += 1024) {
        const uint2 r = ranges[i];
        const uint32_t n = r.y - r.x;
        order[atomicAdd(&start[n ? 32 - __clz(n) : 0], 1u)] = (uint32_t)i;
    }
}

int vcr_launch_tile_ranges(BinningState& b, ImageState& img, int R, int tiles, cudaStream_t s) {
    { VcrProfScope m(VCR_K_MEMSET, s); VCR_CUDA_TRY(cudaMemsetAsync(img.ranges, 0, (size_t)tiles * sizeof(uint2), s)); }
    if (R > 0) {
        VcrProfScope prof(VCR_K_TILE_RANGES, s);
        tile_ranges_kernel<<<((uint32_t)R + 1023u) / 1024u, 256, 0, s>>>((uint32_t)R, b.tile_keys, img.ranges);
        VCR_LAUNCH_CHECK("tile_ranges_kernel");
    }
    {
        VcrProfScope prof(VCR_K_TILE_RANGES, s);
        tile_order_kernel<<<1, 1024, 0, s>>>(img.ranges, tiles, img.tile_order);
        VCR_LAUNCH_CHECK("tile_order_kernel");
    }
    return 0;
}
