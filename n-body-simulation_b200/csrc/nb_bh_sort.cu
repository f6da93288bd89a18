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

cudaError_t bh_radix_sort(BhTree *t, unsigned n, int bits, cudaStream_t s, unsigned *launches,
                          unsigned offset)
{
    if (n == 0) return cudaSuccess;
    const unsigned nb = bh_sort_blocks(n);
    int passes = (bits + 7) / 8;
    if (passes & 1) ++passes; // an even number of passes leaves the result in buffer 0
    for (int p = 0; p < passes; ++p) {
        const int in = p & 1, out = in ^ 1, shift = 8 * p;
        digit_histogram<<<nb, SORT_THREADS, 0, s>>>(t->key[in] + offset, n, shift, nb, t->hist);
        scan_rows<<<256, 1024, 0, s>>>(t->hist, nb, t->hist + (size_t)256 * nb);
        scatter<<<nb, SORT_THREADS, 0, s>>>(t->key[in] + offset, t->val[in] + offset, t->key[out] + offset,
                                            t->val[out] + offset, n, shift, nb, t->hist,
                                            t->hist + (size_t)256 * nb);
        if (launches) *launches += 3;
    }
    return cudaGetLastError();
}

// ---- bottom-up merge of sorted runs (multi-rank engines) --------------------------------------
// Every rank sorts the keys of its own slice; the sorted (key, index) runs are all-gathered and
// merged here: one pass merges adjacent runs of length `seg` pairwise (merge path: the CTA of an
// output tile finds its two input ranges by binary search on the tile's diagonals, stages them in
// shared memory, and every thread merges 8 outputs from its own diagonal).  Ties take the left
// run first, so equal keys stay in index order like in the stable radix sort: the merged
// sequence IS what bh_radix_sort gives on all keys.  ceil(log2(W)) passes of 24 B/key instead of
// 8 radix passes of 32 B/key over all bodies on every rank.
namespace {
constexpr int MERGE_THREADS = 256;
constexpr int MERGE_ITEMS = 8;
constexpr int MERGE_TILE = MERGE_THREADS * MERGE_ITEMS;

// how many elements of A are among the first d outputs of merge(A, B); ties: A first
__device__ __forceinline__ unsigned merge_path(const uint64_t *A, unsigned la, const uint64_t *B,
                                               unsigned lb, unsigned d)
{
    unsigned lo = d > lb ? d - lb : 0u, hi = d < la ? d : la;
    while (lo < hi) {
        const unsigned mid = (lo + hi) >> 1;
        if (A[mid] <= B[d - 1 - mid]) lo = mid + 1;
        else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(MERGE_THREADS)
    merge_pass_kernel(const uint64_t *__restrict__ kin, const unsigned *__restrict__ vin,
                      uint64_t *__restrict__ kout, unsigned *__restrict__ vout, unsigned total,
                      unsigned seg, unsigned tiles_per_pair)
{
    __shared__ uint64_t sk[MERGE_TILE];
    __shared__ unsigned sv[MERGE_TILE];
    __shared__ unsigned s_a[2];
    const unsigned pair = blockIdx.x / tiles_per_pair, tile = blockIdx.x % tiles_per_pair;
    const unsigned long long p0l = (unsigned long long)pair * 2ull * seg;
    if (p0l >= total) return;
    const unsigned p0 = (unsigned)p0l;
    const unsigned la = min(seg, total - p0), lb = min(seg, total - p0 - la);
    const unsigned len = la + lb, d0 = tile * MERGE_TILE;
    if (d0 >= len) return;
    const unsigned d1 = min(d0 + (unsigned)MERGE_TILE, len);
    const uint64_t *A = kin + p0, *B = kin + p0 + la;
    if (threadIdx.x == 0) s_a[0] = merge_path(A, la, B, lb, d0);
    if (threadIdx.x == 32) s_a[1] = merge_path(A, la, B, lb, d1);
    __syncthreads();
    const unsigned a0 = s_a[0], a1 = s_a[1], b0 = d0 - a0;
    const unsigned na = a1 - a0, nb = (d1 - a1) - b0, cnt = na + nb;
    for (unsigned i = threadIdx.x; i < cnt; i += MERGE_THREADS) {
        const unsigned src = i < na ? p0 + a0 + i : p0 + la + b0 + (i - na);
        sk[i] = kin[src];
        sv[i] = vin[src];
    }
    __syncthreads();
    const unsigned t0 = min(threadIdx.x * (unsigned)MERGE_ITEMS, cnt), t1 = min(t0 + (unsigned)MERGE_ITEMS, cnt);
    unsigned sa = merge_path(sk, na, sk + na, nb, t0), sb = t0 - sa;
    for (unsigned o = t0; o < t1; ++o) {
        const bool take_a = sb >= nb || (sa < na && sk[sa] <= sk[na + sb]);
        const unsigned idx = take_a ? sa++ : na + sb++;
        kout[p0 + d0 + o] = sk[idx];
        vout[p0 + d0 + o] = sv[idx];
    }
}
} // namespace

cudaError_t bh_merge_runs(BhTree *t, unsigned total, unsigned run, cudaStream_t s, unsigned *launches)
{
    int in = 0;
    for (unsigned long long seg = run; seg < total; seg *= 2) {
        const unsigned tpp = (unsigned)((2ull * seg + MERGE_TILE - 1) / MERGE_TILE);
        const unsigned pairs = (unsigned)((total + 2ull * seg - 1) / (2ull * seg));
        merge_pass_kernel<<<pairs * tpp, MERGE_THREADS, 0, s>>>(t->key[in], t->val[in], t->key[in ^ 1],
                                                                 t->val[in ^ 1], total, (unsigned)seg, tpp);
        in ^= 1;
        if (launches) ++*launches;
    }
    t->sorted = in;
    return cudaGetLastError();
}

} // namespace nb
