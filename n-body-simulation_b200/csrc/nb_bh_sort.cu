// nb_bh_sort.cu -- stable LSD radix sort of (Morton key, body index) pairs,
// 8-bit digits, hand-written for sm_100a (no CUB on the product path).
//
// Replaces sort_ents of the reference (cuda/barnes-hut.cu:459-488), which walks
// the finished tree to put bodies in depth-first order; here the order comes
// first and the tree is built from it.
//
// One pass = three launches:
//   digit_histogram : per 2048-key tile, 256 bin counts -> hist[bin][tile]
//   exclusive_scan  : over hist in (bin, tile) order -> global base of every (bin, tile)
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

// Exclusive prefix sum of `len` counters in place, one CTA of 1024 threads: each
// thread owns a contiguous chunk, the chunk totals are scanned in shared memory.
__global__ void __launch_bounds__(1024) exclusive_scan(unsigned *__restrict__ a, unsigned len)
{
    __shared__ unsigned part[1024];
    const unsigned chunk = (len + 1023u) / 1024u;
    const unsigned b = threadIdx.x * chunk;
    const unsigned e = min(b + chunk, len);
    unsigned s = 0;
    for (unsigned i = b; i < e; ++i) s += a[i];
    part[threadIdx.x] = s;
    __syncthreads();
    // Hillis-Steele over 1024 partials
    for (unsigned d = 1; d < 1024; d <<= 1) {
        const unsigned v = threadIdx.x >= d ? part[threadIdx.x - d] : 0u;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned run = part[threadIdx.x] - s; // exclusive base of this chunk
    for (unsigned i = b; i < e; ++i) {
        const unsigned v = a[i];
        a[i] = run;
        run += v;
    }
}

__global__ void __launch_bounds__(SORT_THREADS)
    scatter(const uint64_t *__restrict__ kin, const unsigned *__restrict__ vin,
            uint64_t *__restrict__ kout, unsigned *__restrict__ vout, unsigned n, int shift,
            unsigned nblocks, const unsigned *__restrict__ base)
{
    __shared__ unsigned wcount[SORT_WARPS][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    for (int i = threadIdx.x; i < SORT_WARPS * 256; i += SORT_THREADS) (&wcount[0][0])[i] = 0;
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
        unsigned run = base[(size_t)t * nblocks + blockIdx.x];
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
        exclusive_scan<<<1, 1024, 0, s>>>(t->hist, 256u * nb);
        scatter<<<nb, SORT_THREADS, 0, s>>>(t->key[in], t->val[in], t->key[out], t->val[out], n,
                                            shift, nb, t->hist);
        if (launches) *launches += 3;
    }
    return cudaGetLastError();
}

} // namespace nb
