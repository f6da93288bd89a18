// nb_bh_sort.cu -- stable LSD radix sort of (Morton key, body index) pairs,
// 8-bit digits, hand-written for sm_100a (no CUB on the product path).
//
// Replaces sort_ents of the reference (cuda/barnes-hut.cu:459-488), which walks
// the finished tree to put bodies in depth-first order; here the order comes
// first and the tree is built from it.
//
// One pass = three launches:
//   digit_histogram : per 2048-key tile, 256 bin counts -> hist[bin][tile]
//   scan_rows       : CTA d scans row d of hist (tile order) and reports the row total;
//                     the 256 totals are scanned inside scatter -> base of every (bin, tile)
//   scatter         : re-reads the tile; a warp owns 256 consecutive keys and ranks
//                     them 32 at a time with __match_any_sync, so equal digits keep
//                     their input order (tile, warp, round, lane) = stable.
// Traffic per pass: keys read twice, values once, both written once = 32 B/key;
// HBM-bound in principle, launch-bound at 1M keys (24 launches of a few us).
#include "nb_bh.cuh"

namespace nb {
namespace {

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 8;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS; // 2048 keys per CTA
constexpr int SORT_WARPS = SORT_THREADS / 32;

__global__ void __launch_bounds__(SORT_THREADS)
    digit_histogram(const uint64_t *__restrict__ keys, unsigned n, int shift, unsigned nblocks,
                    unsigned *__restrict__ hist)
{
    __shared__ unsigned h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const unsigned base = blockIdx.x * SORT_TILE;
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r) {
        const unsigned i = base + r * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(unsigned)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// Row d of hist (the counts of digit d in every tile, tile order) is scanned in
// place by CTA d; the row total goes to totals[d].  The offset of the row itself
// (all smaller digits) is added by the scatter kernel from the 256 totals.
__global__ void __launch_bounds__(1024)
    scan_rows(unsigned *__restrict__ hist, unsigned nblocks, unsigned *__restrict__ totals)
{
    __shared__ unsigned wsum[32];
    __shared__ unsigned carry;
    unsigned *row = hist + (size_t)blockIdx.x * nblocks;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (unsigned b0 = 0; b0 < nblocks; b0 += 1024) {
        const unsigned i = b0 + threadIdx.x;
        const unsigned v = i < nblocks ? row[i] : 0u;
        unsigned x = v; // inclusive warp scan
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) wsum[warp] = x;
        __syncthreads();
        if (warp == 0) {
            unsigned w = wsum[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            wsum[lane] = w; // inclusive over warps
        }
        __syncthreads();
        const unsigned before = carry + (warp ? wsum[warp - 1] : 0u);
        if (i < nblocks) row[i] = before + x - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}

__global__ void __launch_bounds__(SORT_THREADS)
    scatter(const uint64_t *__restrict__ kin, const unsigned *__restrict__ vin,
            uint64_t *__restrict__ kout, unsigned *__restrict__ vout, unsigned n, int shift,
            unsigned nblocks, const unsigned *__restrict__ base,
            const unsigned *__restrict__ totals)
{
    __shared__ unsigned wcount[SORT_WARPS][256];
    __shared__ unsigned dbase[256]; // keys with a smaller digit, all tiles
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    for (int i = threadIdx.x; i < SORT_WARPS * 256; i += SORT_THREADS) (&wcount[0][0])[i] = 0;
    { // exclusive scan of the 256 digit totals (8 warps x 32 digits)
        const unsigned t = totals[threadIdx.x];
        unsigned x = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        dbase[threadIdx.x] = x - t; // exclusive inside the warp's 32 digits
        __syncthreads();
        unsigned before = 0;
        for (int w = 0; w < warp; ++w) before += dbase[w * 32 + 31] + totals[w * 32 + 31];
        __syncthreads();
        dbase[threadIdx.x] = before + x - t;
    }
    __syncthreads();

    uint64_t key[SORT_ITEMS];
    unsigned val[SORT_ITEMS], loc[SORT_ITEMS];
    const unsigned wbase = blockIdx.x * SORT_TILE + warp * (32 * SORT_ITEMS);
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r) {
        const unsigned i = wbase + r * 32 + lane;
        const bool valid = i < n;
        key[r] = valid ? kin[i] : 0;
        val[r] = valid ? vin[i] : 0;
        const unsigned d = valid ? ((unsigned)(key[r] >> shift) & 255u) : 256u; // 256: nobody's digit
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const unsigned rank = __popc(peers & lt);
        unsigned off = 0;
        if (valid) off = wcount[warp][d];
        __syncwarp();
        if (valid && rank == 0) wcount[warp][d] = off + __popc(peers);
        __syncwarp();
        loc[r] = off + rank;
    }
    __syncthreads();
    { // digit t: running offsets over the CTA's warps + the global base of (digit, tile)
        const unsigned t = threadIdx.x;
        unsigned run = dbase[t] + base[(size_t)t * nblocks + blockIdx.x];
#pragma unroll
        for (int w = 0; w < SORT_WARPS; ++w) {
            const unsigned c = wcount[w][t];
            wcount[w][t] = run;
            run += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; ++r) {
        const unsigned i = wbase + r * 32 + lane;
        if (i < n) {
            const unsigned d = (unsigned)(key[r] >> shift) & 255u;
            const unsigned pos = wcount[warp][d] + loc[r];
            kout[pos] = key[r];
            vout[pos] = val[r];
        }
    }
}

} // namespace

cudaError_t bh_exclusive_scan(unsigned *a, unsigned len, unsigned *total, cudaStream_t s)
{
    scan_rows<<<1, 1024, 0, s>>>(a, len, total);
    return cudaGetLastError();
}

unsigned bh_sort_blocks(unsigned n) { return n ? (n + SORT_TILE - 1) / SORT_TILE : 1u; }

cudaError_t bh_radix_sort(BhTree *t, unsigned n, int bits, cudaStream_t s, unsigned *launches)
{
    if (n == 0) return cudaSuccess;
    const unsigned nb = bh_sort_blocks(n);
    int passes = (bits + 7) / 8;
    if (passes & 1) ++passes; // an even number of passes leaves the result in buffer 0
    for (int p = 0; p < passes; ++p) {
        const int in = p & 1, out = in ^ 1, shift = 8 * p;
        digit_histogram<<<nb, SORT_THREADS, 0, s>>>(t->key[in], n, shift, nb, t->hist);
        scan_rows<<<256, 1024, 0, s>>>(t->hist, nb, t->hist + (size_t)256 * nb);
        scatter<<<nb, SORT_THREADS, 0, s>>>(t->key[in], t->val[in], t->key[out], t->val[out], n,
                                            shift, nb, t->hist, t->hist + (size_t)256 * nb);
        if (launches) *launches += 3;
    }
    return cudaGetLastError();
}

} // namespace nb
