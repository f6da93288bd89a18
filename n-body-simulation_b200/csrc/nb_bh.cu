// nb_bh.cu -- Barnes-Hut on sm_100a: bounding cube, Morton keys, level-wise
// octree build from the sorted keys, bottom-up centre of mass, warp-cooperative
// THETA traversal with the leapfrog epilogue fused in.  See nb_bh.cuh for the
// correspondence with the reference (serial/barnes-hut.c, cuda/barnes-hut.cu).
#include "nb_bh.cuh"
#include <stdlib.h>
#include "nb_common.cuh"

namespace nb {
namespace {

constexpr int BH_THREADS = 256;

// ---------------------------------------------------------------- bbox ----
// get_bounding_box (serial/barnes-hut.c:326-338; cuda/barnes-hut.cu:307-349):
// max over bodies of max(|x|,|y|,|z|); the cube side is twice that.  |v| >= 0,
// so the IEEE bit pattern orders like the value and atomicMax works on it.
__global__ void __launch_bounds__(BH_THREADS)
    bbox_kernel(const double4 *__restrict__ q, unsigned n, unsigned long long *out)
{
    double m = 0.0;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double4 b = q[i];
        m = fmax(m, fmax(fabs(b.x), fmax(fabs(b.y), fabs(b.z))));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.0)
        atomicMax(out, (unsigned long long)__double_as_longlong(m));
}

// ---------------------------------------------------------------- keys ----
// The octant path of get_indx_loc (serial/barnes-hut.c:114-133), followed
// BH_MAX_LEVEL levels down with the reference's own arithmetic (centre +=
// +-border/4, border /= 2, `>=` goes high), so a body lands in the same cell
// as in the reference even where the rounded centre decides.  Digit of level 1
// in the top bits: sorting the keys sorts the bodies depth-first.
__global__ void __launch_bounds__(BH_THREADS)
    morton_kernel(const double4 *__restrict__ q, unsigned n, const double *__restrict__ bbox,
                  uint64_t *__restrict__ key, unsigned *__restrict__ val)
{
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double4 b = q[i];
    double border = 2.0 * bbox[0];
    double cx = 0.0, cy = 0.0, cz = 0.0;
    uint64_t k = 0;
#pragma unroll 1
    for (int l = 0; l < BH_MAX_LEVEL; ++l) {
        const double q4 = border / 4;
        const int z = b.z >= cz, y = b.y >= cy, x = b.x >= cx;
        k = (k << 3) | (uint64_t)(z * 4 + y * 2 + x);
        cx += x ? q4 : -q4;
        cy += y ? q4 : -q4;
        cz += z ? q4 : -q4;
        border /= 2;
    }
    key[i] = k;
    val[i] = i;
}

// multi-rank: keys of the bodies [begin, begin + count) only, written in place; indices past the
// last real body get the all-ones key (63-bit keys never reach it), so they sort behind everybody
__global__ void __launch_bounds__(BH_THREADS)
    morton_slice_kernel(const double4 *__restrict__ q, unsigned n, unsigned begin, unsigned count,
                        const double *__restrict__ bbox, uint64_t *__restrict__ key,
                        unsigned *__restrict__ val)
{
    const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count) return;
    const unsigned i = begin + t;
    uint64_t k = ~0ull;
    if (i < n) {
        const double4 b = q[i];
        double border = 2.0 * bbox[0];
        double cx = 0.0, cy = 0.0, cz = 0.0;
        k = 0;
#pragma unroll 1
        for (int l = 0; l < BH_MAX_LEVEL; ++l) {
            const double q4 = border / 4;
            const int z = b.z >= cz, y = b.y >= cy, x = b.x >= cx;
            k = (k << 3) | (uint64_t)(z * 4 + y * 2 + x);
            cx += x ? q4 : -q4;
            cy += y ? q4 : -q4;
            cz += z ? q4 : -q4;
            border /= 2;
        }
    }
    key[i] = k;
    val[i] = i;
}

// the cube of all bodies from the ranks' partial maxima (max is exact: any order gives the
// single-rank value)
__global__ void bbox_combine_kernel(const double *__restrict__ parts, unsigned world, double *bbox)
{
    if (threadIdx.x == 0) {
        double m = 0.0;
        for (unsigned r = 0; r < world; ++r) m = fmax(m, parts[r]);
        bbox[0] = m;
    }
}

// positions in Morton order + the list of sorted indices this rank integrates
__global__ void __launch_bounds__(BH_THREADS)
    gather_sorted_kernel(const double4 *__restrict__ q, const unsigned *__restrict__ perm,
                         unsigned n, double4 *__restrict__ spos)
{
    const unsigned k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) spos[k] = q[perm[k]];
}

// Order-preserving compaction of the sorted indices whose body lies in
// [own_begin, own_end) (multi-rank runs: the bodies this rank integrates, in
// Morton order): per-tile counts, an exclusive scan of the counts
// (bh_exclusive_scan), then every tile writes its members behind its base.
constexpr int OWN_TILE = 1024;

__device__ __forceinline__ unsigned own_rank_in_tile(bool mine, unsigned *wsum, unsigned *total)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned bal = __ballot_sync(0xffffffffu, mine);
    if (lane == 0) wsum[warp] = __popc(bal);
    __syncthreads();
    unsigned before = 0, all = 0;
    for (int w = 0; w < OWN_TILE / 32; ++w) {
        const unsigned c = wsum[w];
        before += w < warp ? c : 0u;
        all += c;
    }
    *total = all;
    return before + __popc(bal & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(OWN_TILE)
    own_count_kernel(const unsigned *__restrict__ perm, unsigned n, unsigned own_begin,
                     unsigned own_end, unsigned *__restrict__ counts)
{
    __shared__ unsigned wsum[OWN_TILE / 32];
    const unsigned k = blockIdx.x * OWN_TILE + threadIdx.x;
    const bool mine = k < n && perm[k] >= own_begin && perm[k] < own_end;
    unsigned total;
    own_rank_in_tile(mine, wsum, &total);
    if (threadIdx.x == 0) counts[blockIdx.x] = total;
}

__global__ void __launch_bounds__(OWN_TILE)
    own_write_kernel(const unsigned *__restrict__ perm, unsigned n, unsigned own_begin,
                     unsigned own_end, const unsigned *__restrict__ base, unsigned *__restrict__ own)
{
    __shared__ unsigned wsum[OWN_TILE / 32];
    const unsigned k = blockIdx.x * OWN_TILE + threadIdx.x;
    const bool mine = k < n && perm[k] >= own_begin && perm[k] < own_end;
    unsigned total;
    const unsigned r = own_rank_in_tile(mine, wsum, &total);
    if (mine) own[base[blockIdx.x] + r] = k;
}

// --------------------------------------------------------------- build ----
// Nodes are numbered level by level: level L occupies [base_L, base_L + cnt_L),
// base_L = sum of the counts above.  The kernel of level L splits every node of
// that level by the key digit of the level into <= 8 children -- a child with
// one body is a leaf (the body), with >= 2 bodies a new node of level L+1 (the
// reference's "branch iff the cell holds >= 2 bodies", serial/barnes-hut.c:
// 190-246).  Eight lanes work on one node, one octant each.
struct BuildArgs {
    const uint64_t *key;
    unsigned *first, *count;
    unsigned char *level;
    int *child;
    unsigned *level_count;
    unsigned *flags;
    unsigned node_cap;
    const double4 *spos;
    ChildRec *cblock;
    unsigned *pslot;
};

__device__ __forceinline__ unsigned level_base(const unsigned *level_count, int L)
{
    unsigned b = 0;
    for (int l = 0; l < L; ++l) b += level_count[l];
    return b;
}
// nodes of level L that exist in the pool (an overflowing level is cut, and flagged)
__device__ __forceinline__ unsigned level_nodes(const unsigned *level_count, int L, unsigned base,
                                                unsigned cap)
{
    const unsigned c = level_count[L];
    return base >= cap ? 0u : min(c, cap - base);
}

__global__ void __launch_bounds__(BH_THREADS) build_level_kernel(BuildArgs a, int L)
{
    const unsigned base = level_base(a.level_count, L);
    const unsigned cnt = level_nodes(a.level_count, L, base, a.node_cap);
    const unsigned next_base = base + a.level_count[L];
    const int oct = threadIdx.x & 7;
    const unsigned gmask = 0xffu << ((threadIdx.x & 31) & ~7); // this node's 8 lanes
    const int shift = 3 * (BH_MAX_LEVEL - 1 - L);
    // Node ids of the next level come from one counter.  One atomicAdd per node made the
    // populous levels atomic-bound (3M same-address atomics at 16M bodies); the CTA's 32 nodes
    // are scanned in shared memory and take their ids with ONE atomicAdd.  Every thread of the
    // CTA runs the same trip count (idle groups vote with zero children).
    constexpr unsigned GPC = BH_THREADS / 8; // nodes per CTA and trip
    __shared__ unsigned s_off[GPC];
    __shared__ unsigned s_base;
    const unsigned grp = threadIdx.x >> 3;
    for (unsigned g0 = blockIdx.x * GPC; g0 < cnt; g0 += gridDim.x * GPC) {
        const unsigned g = g0 + grp;
        const bool live = g < cnt;
        const unsigned nd = base + (live ? g : 0u);
        const unsigned f = live ? a.first[nd] : 0u, c = live ? a.count[nd] : 0u;
        // lower bound of digit `oct` inside [f, f + c)
        unsigned lo = f, hi = f + c;
        while (lo < hi) {
            const unsigned mid = (lo + hi) >> 1;
            if ((unsigned)((a.key[mid] >> shift) & 7u) < (unsigned)oct) lo = mid + 1;
            else hi = mid;
        }
        const unsigned start = lo;
        unsigned end = __shfl_down_sync(gmask, start, 1, 8);
        if (oct == 7) end = f + c;
        const unsigned k = end - start;
        const bool branch = k >= 2;
        const unsigned bal = (__ballot_sync(gmask, branch) >> ((threadIdx.x & 31) & ~7)) & 0xffu;
        if (oct == 0) s_off[grp] = (unsigned)__popc(bal);
        __syncthreads();
        if (threadIdx.x < 32) {
            const unsigned v = s_off[threadIdx.x];
            unsigned x = v; // inclusive scan over the CTA's 32 nodes
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned y = __shfl_up_sync(0xffffffffu, x, o);
                if ((int)threadIdx.x >= o) x += y;
            }
            if (threadIdx.x == 31 && x) s_base = atomicAdd(&a.level_count[L + 1], x);
            s_off[threadIdx.x] = x - v;
        }
        __syncthreads();
        const unsigned slot = s_base + s_off[grp]; // s_base is only read when the CTA made nodes
        __syncthreads(); // the next trip overwrites s_off / s_base
        if (!live) continue;
        int ch = -1;
        if (k == 1) ch = (int)start;
        else if (branch) {
            const unsigned id = next_base + slot + __popc(bal & ((1u << oct) - 1u));
            if (id < a.node_cap) {
                a.first[id] = start;
                a.count[id] = k;
                a.level[id] = (unsigned char)(L + 1);
                ch = -(int)id - 2;
            } else {
                a.flags[0] = 1; // node pool exhausted (the reference exits: openmp/barnes-hut.c:158-161)
            }
        }
        a.child[(size_t)nd * 8 + oct] = ch;
        // the traversal's view: non-empty children packed to the front of the node's block
        const unsigned nonempty = (__ballot_sync(gmask, k != 0) >> ((threadIdx.x & 31) & ~7)) & 0xffu;
        const unsigned at = nd * 8u + (unsigned)__popc(nonempty & ((1u << oct) - 1u));
        if (k == 1) {
            ChildRec r;
            r.c = a.spos[start];
            r.b2 = -1.0;
            r.ent = (int)start;
            r.code = 0;
            a.cblock[at] = r;
        } else if (ch <= -2) {
            // centre of mass and border^2 follow in com_level_kernel, the child count when the
            // child itself is split (next level)
            a.cblock[at].ent = ch;
            a.cblock[at].code = (int)((unsigned)(-ch - 2) | (L + 1 == BH_MAX_LEVEL ? BH_CODE_BUCKET : 0u));
            a.pslot[-ch - 2] = at;
        } else if (k >= 2) {
            // node pool exhausted (flagged, the result is refused): a massless stand-in keeps
            // the walk inside valid memory
            ChildRec r;
            r.c = a.spos[start];
            r.c.w = 0.0;
            r.b2 = -1.0;
            r.ent = (int)start;
            r.code = 0;
            a.cblock[at] = r;
        }
        if (oct == 0) { // how many records the walk reads when it opens this node
            const unsigned code = nd | ((unsigned)(__popc(nonempty) - 1) << BH_CODE_NC_SHIFT);
            if (nd == 0) a.flags[3] = code;
            else a.cblock[a.pslot[nd]].code = (int)code;
        }
    }
}

// The reference accepts a cell when border / dist < THETA (serial/barnes-hut.c:405).  With
// r2 = dist^2 + eps2 -- which the interaction needs anyway -- this is
//     border^2 / THETA^2 + eps2 < r2,
// one FP64 compare per candidate instead of add + multiply + compare.  The left side is a
// property of the cell's level and is stored in its record.  THETA = 0 gives +inf (never
// accept: the direct sum); bodies carry -1 (always accept).  Like the squared form it can differ
// from the reference's divide-and-compare only when border/dist is within rounding of THETA.
__host__ __device__ __forceinline__ double bh_threshold(double border, double theta, double eps2)
{
    return (border * border) / (theta * theta) + eps2;
}

// ----------------------------------------------------------------- COM ----
// center_of_mass (serial/barnes-hut.c:285-317): children folded in octant
// order with the reference's incremental weighted mean, written as it writes it
// (x*m/nm + c*cm/nm), one thread per node, deepest level first.  A level-21
// node whose bodies share one key ("bucket") folds its bodies in sorted order.
__global__ void __launch_bounds__(BH_THREADS)
    com_level_kernel(const double4 *__restrict__ spos, const int *__restrict__ child,
                     const unsigned *__restrict__ first, const unsigned *__restrict__ count,
                     const unsigned *__restrict__ level_count, double4 *__restrict__ com, int L,
                     unsigned node_cap, const double *__restrict__ bbox,
                     ChildRec *__restrict__ cblock, const unsigned *__restrict__ pslot,
                     double theta, double eps2)
{
    // border^2 at depth L: (max / 2^L)^2, the halvings are exact.  The record carries the
    // acceptance threshold in the form the walk tests it (bh_threshold).
    double border = 2.0 * bbox[0];
    for (int l = 0; l < L; ++l) border /= 2;
    const double b2 = bh_threshold(border, theta, eps2);
    const unsigned base = level_base(level_count, L);
    const unsigned cnt = level_nodes(level_count, L, base, node_cap);
    for (unsigned g = blockIdx.x * blockDim.x + threadIdx.x; g < cnt; g += gridDim.x * blockDim.x) {
        const unsigned nd = base + g;
        double cx = 0, cy = 0, cz = 0, cm = 0;
        auto fold = [&](const double4 &b) {
            const double nm = __dadd_rn(cm, b.w);
            cx = __dadd_rn(__ddiv_rn(__dmul_rn(b.x, b.w), nm), __ddiv_rn(__dmul_rn(cx, cm), nm));
            cy = __dadd_rn(__ddiv_rn(__dmul_rn(b.y, b.w), nm), __ddiv_rn(__dmul_rn(cy, cm), nm));
            cz = __dadd_rn(__ddiv_rn(__dmul_rn(b.z, b.w), nm), __ddiv_rn(__dmul_rn(cz, cm), nm));
            cm = nm;
        };
        if (L == BH_MAX_LEVEL) {
            const unsigned f = first[nd], c = count[nd];
            for (unsigned k = 0; k < c; ++k) fold(spos[f + k]);
        } else {
#pragma unroll 1
            for (int o = 0; o < 8; ++o) {
                const int ch = child[(size_t)nd * 8 + o];
                if (ch == -1) continue;
                fold(ch >= 0 ? spos[ch] : com[-ch - 2]);
            }
        }
        com[nd] = make_double4(cx, cy, cz, cm);
        if (L > 0) {
            ChildRec *r = &cblock[pslot[nd]];
            r->c = make_double4(cx, cy, cz, cm);
            r->b2 = b2;
        }
    }
}

// ----------------------------------------------------------- traversal ----
// get_acceleration_rec (serial/barnes-hut.c:389-416) for 32 bodies at a time
// (consecutive in Morton order), like acceleration_w_stack of the reference
// (cuda/barnes-hut.cu:538-629): one depth-first walk per warp over a stack in
// shared memory.  Unlike the reference kernel -- which opens a node for the
// whole warp unless ALL lanes accept it -- every stack entry carries the mask of
// lanes still descending: a lane that accepted a cell (border/dist < THETA)
// drops out of its subtree.  Each body therefore sees exactly the serial
// reference's interaction list, in the reference's order (children 0..7,
// depth first), and agrees with it to rounding, not just within the THETA
// envelope.  An opened cell's child records (one 384-byte block) are fetched by
// the warp in one coalesced copy and tested in one go (one memory round trip per
// OPENED cell instead of one per visited node); only cells that some lane must
// open go through the stack.  The sum order therefore is "a cell's accepted
// children first, then its opened children's subtrees" -- the reference's terms,
// regrouped.  The loop over a cell's children is straight-line code (no vote, no
// branch per child: ncu showed the walk bound by issue slots, 45 instructions per
// child of which 19 FP64): every lane computes the interaction and adds it under
// a select, and records in a bit mask what it did not take; the votes happen once
// per cell, only for the children that some lane must open.
//
// Acceptance is tested as border^2 < THETA^2 * dist^2 -- the reference's
// border / dist < THETA (:405) without the FP64 divide and square root; the two
// differ only when border/dist is within an ulp of THETA.
constexpr int TRAV_WARPS = 4;
constexpr int TRAV_STACK = 7 * (BH_MAX_LEVEL + 1) + 16; // <= 7 siblings pending per level + 8

struct TravArgs {
    const double4 *spos, *com;
    const ChildRec *cblock;
    const unsigned *first, *count, *perm, *own;
    const unsigned *cta_order; // nullptr: CTAs in body order
    unsigned *cta_cost;
    const double *bbox;
    unsigned *flags;
    unsigned long long *counters;
    unsigned n_own_static; // used when own == nullptr (all bodies, identity list)
    BhParams p;
};

// m * (r2)^(-3/2), the 7-operation cube of nb_direct_f64.cu
__device__ __forceinline__ double bh_inv_r3_times_m(double r2, double m)
{
    const double y0 = rsqrt_approx_f64(r2);
    const double y02 = y0 * y0;
    const double e = fma(-r2, y02, 1.0);
    const double pe = fma(1.875, e, 1.5);
    const double q = e * pe;
    const double my = m * y0;
    const double u = y02 * my;
    return fma(u, q, u);
}

#ifndef NB_BH_TRAV_MINB
#define NB_BH_TRAV_MINB 8 // 64 registers, 32 warps per SM: no spills, constants stay in registers; 10 (48 regs) measured 7 % slower
#endif
// SELF: test "is this record the lane's own body" (:411).  With eps2 > 0 the test changes nothing:
// the own body sits at distance 0, so its term is fma(0, m / eps2^1.5, a) = a exactly, as if it had
// been skipped -- the walk then runs without the compare (and without a stack trip for the own
// leaf).  eps2 below 1e-60 (0 * inf) and the counting walk keep it.
// PREFETCH: when a cell is pushed, three lanes ask L2 for the three lines of its child block.  Pays
// when the tree is far larger than L2 or a rank walks scattered pieces of it (misses stall a walk
// that has 8 warps per sub-partition to hide them); costs 3 issue slots per push otherwise.
template <bool COUNT, bool SELF, bool PREFETCH>
__global__ void __launch_bounds__(TRAV_WARPS * 32, NB_BH_TRAV_MINB) traverse_kernel(const TravArgs a)
{
    __shared__ uint2 s_stack[TRAV_WARPS][TRAV_STACK]; // {cell id | depth << 27, lanes that must open it}
    __shared__ ChildRec s_child[TRAV_WARPS][2][8];

    // pointers and scalars in registers: the walk below is issue-bound, not FP64-bound
    const double4 *__restrict__ spos = a.spos;
    const double4 *__restrict__ com = a.com;
    const ChildRec *__restrict__ cblock = a.cblock;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned n_own = a.own ? a.flags[2] : a.n_own_static;
    const double eps2 = a.p.eps2;
    uint2 *stk = s_stack[warp];
    int flip = 0;
    // One CTA per 128 bodies.  NB_BH_TRAV_DYNAMIC (measured, not shipped) makes the grid resident
    // and lets every warp fetch its next 32 bodies from a counter, against the idea that CTAs wait
    // for their slowest warp and the grid drains for a whole walk at the end: 1M Plummer 4.84 ->
    // 5.05 ms, 16M 82.0 -> 83.4 ms -- the hardware's block scheduler already does it better.
    for (;;) {
    unsigned wi = 0, cta_id = 0;
#ifndef NB_BH_TRAV_DYNAMIC
    {
        // CTAs start heaviest first (cta_order_kernel): the grid then ends on its lightest walks
        const unsigned n_ctas = (n_own + TRAV_WARPS * 32 - 1) / (TRAV_WARPS * 32);
        if (blockIdx.x >= n_ctas) break;
        cta_id = a.cta_order ? a.cta_order[blockIdx.x] : blockIdx.x;
        wi = cta_id * TRAV_WARPS + warp;
    }
#else
    if (lane == 0) wi = atomicAdd(&a.flags[4], 1u);
    wi = __shfl_sync(0xffffffffu, wi, 0);
#endif
    if ((unsigned long long)wi * 32ull >= n_own) break;
    const unsigned j = wi * 32 + lane;
    const bool valid = j < n_own;
    const unsigned k = valid ? (a.own ? a.own[j] : j) : 0u;
    const double4 me = spos[k];
    double ax = 0, ay = 0, az = 0;
    unsigned n_inter = 0, n_visit = 0;
    unsigned n_tests = 0; // records this WARP tested (17 FP64 warp instructions each), same in every lane
    int sp = 0;
    const unsigned all = __ballot_sync(0xffffffffu, valid);

    // One candidate (a body or a cell) against the warp's 32 bodies.  A lane takes it if it is
    // one of the lanes the candidate is meant for (`mine`), it is far enough (border^2 <
    // THETA^2 dist^2; bodies carry border^2 = -1) and it is not the lane's own body (:411).
    // Returns whether the lane took it; for a cell the other lanes it was meant for must open
    // it.  Straight-line code: the interaction is computed by every lane and added under a
    // predicate -- the walk is bound by issue slots, and a vote plus a branch per candidate
    // cost more than the few candidates that no lane takes.
    auto consider = [&](const double4 &c, double thr, int ent, bool mine) -> bool {
        const double dx = c.x - me.x, dy = c.y - me.y, dz = c.z - me.z;
        // dist^2 + eps2 in one FMA chain: within rounding of the reference's mul/add form
        // (:101-104, :172), which moves an accept/open decision only at the very edge of THETA
        const double r2 = fma(dz, dz, fma(dy, dy, fma(dx, dx, eps2)));
#ifdef NB_BH_INT_THRESHOLD
        // thr < r2 on the bit patterns: r2 > 0 and finite, thr is -1 (bodies), a positive number,
        // +inf (THETA = 0) or NaN (a cube of size 0: never accept) -- as signed 64-bit integers
        // these order exactly like the doubles (NaN's pattern is above +inf: "not below", as for the
        // doubles), and the compare leaves the FP64 pipe
        const bool far = __double_as_longlong(thr) < __double_as_longlong(r2);
#else
        const bool far = thr < r2;
#endif
        const bool take = SELF ? (mine & far & (ent != (int)k)) : (mine & far); // no short circuit
        const double s0 = bh_inv_r3_times_m(r2, c.w);
#ifdef NB_BH_SELECT_S
        const double s = take ? s0 : 0.0; // selected after the fact: a NaN/inf of a lane that does not take it is dropped
        ax = fma(dx, s, ax); ay = fma(dy, s, ay); az = fma(dz, s, az);
#else
        // three predicated DFMAs: the same issue slots as unpredicated ones, and no two FSELs to
        // zero s first (a NaN/inf of a lane that does not take the candidate never reaches the sums)
        // (volatile asm: the compiler would otherwise speculate the FMAs and select the sums,
        // six FSELs; ptxas turns the short branch into predicates)
        if (take) {
            asm volatile("fma.rn.f64 %0, %3, %6, %0;\n\t"
                         "fma.rn.f64 %1, %4, %6, %1;\n\t"
                         "fma.rn.f64 %2, %5, %6, %2;"
                         : "+d"(ax), "+d"(ay), "+d"(az)
                         : "d"(dx), "d"(dy), "d"(dz), "d"(s0));
        }
#endif
        if (COUNT) {
            n_inter += take;
            n_visit += ent >= 0 ? take : mine; // the body's own leaf is never visited
        }
        return take;
    };

    if (all && a.p.n >= 2) { // a single body feels nothing (the root needs >= 2 bodies)
        const double root_border = 2.0 * a.bbox[0]; // = max (serial/barnes-hut.c:470)
        // the root is tested like any cell (:405, :434): same threshold form as the child
        // records carry (border^2 / THETA^2 + eps2 against dist^2 + eps2; THETA = 0 -> +inf)
        const double root_thr = bh_threshold(root_border, a.p.theta, eps2);
        const bool took = consider(com[0], root_thr, -2, valid);
        ++n_tests;
        const unsigned O = __ballot_sync(0xffffffffu, valid && !took);
        if (O) {
            stk[0] = make_uint2(a.flags[3], O);
            sp = 1;
        }
    }
    __syncwarp();
    while (sp > 0) {
        --sp;
        const uint2 top = stk[sp];
        const unsigned nd = top.x & BH_CODE_ID, M = top.y;
        const bool mine = (M >> lane) & 1u;
        if (top.x & BH_CODE_BUCKET) { // bodies sharing one finest cell, summed directly
            const unsigned f = a.first[nd], cn = a.count[nd];
            for (unsigned t = f; t < f + cn; ++t) {
                consider(spos[t], -1.0, (int)t, mine);
                if (COUNT) n_visit -= mine && t != k; // members are not tree nodes
            }
            n_tests += cn;
            continue;
        }
        // The node's child records are one 384-byte block: 24 lanes copy 16 bytes each
        // (one coalesced round trip per opened cell) into shared memory, from where all lanes
        // read the records as broadcasts.  How many records count is in the stack word.  Two
        // buffers alternate, so one barrier per cell orders everything: a lane that runs ahead
        // into the next cell writes the other buffer, and cannot pass that cell's barrier
        // before every lane has left this one.
        const int nc = (int)((top.x >> BH_CODE_NC_SHIFT) & 7u) + 1;
        ChildRec *slot = s_child[warp][flip];
        flip ^= 1;
        if (lane < (int)BH_BLOCK_WORDS)
            reinterpret_cast<uint4 *>(slot)[lane] =
                reinterpret_cast<const uint4 *>(cblock + (size_t)nd * 8)[lane];
        if (sp + nc > TRAV_STACK) { // cannot happen for depth <= BH_MAX_LEVEL; flagged, not silent
            if (lane == 0) a.flags[1] = 1;
            break;
        }
        __syncwarp();
        // Children last to first; a lane notes in `open` (bit c) the candidates it did not take:
        // cells it must descend into (and, once per walk, its own body's leaf).
        n_tests += (unsigned)nc;
        // (one jump per cell into eight straight-line copies was measured 3 % slower than this
        // loop, which the compiler unrolls by four: every copy re-made its constants)
        unsigned open = 0;
        for (int c = nc - 1; c >= 0; --c) {
            const double4 cc = slot[c].c;
            const double2 be = *reinterpret_cast<const double2 *>(&slot[c].b2); // b2 | ent, code
            const bool took = consider(cc, be.x, __double2loint(be.y), mine);
            open = open * 2u + (unsigned)(mine & !took);
        }
        // Cells that some lane must open go to the stack, last child first, so the first child
        // ends on top and is walked first (the reference's depth-first order).  Every lane
        // stores the same words, so each lane's later pop reads them whatever the lanes' order.
        unsigned todo = __reduce_or_sync(0xffffffffu, open);
        while (todo) {
            const int c = 31 - __clz(todo);
            todo &= ~(1u << c);
            const int2 ec = *reinterpret_cast<const int2 *>(&slot[c].ent); // ent, code
            if (ec.x >= 0) continue; // a lane's own body
            if (PREFETCH && lane < 3) {
                const char *blk = reinterpret_cast<const char *>(cblock + (size_t)((unsigned)ec.y & BH_CODE_ID) * 8);
                asm volatile("prefetch.global.L2 [%0];" ::"l"(blk + lane * 128));
            }
            stk[sp++] = make_uint2((unsigned)ec.y, __ballot_sync(0xffffffffu, (open >> c) & 1u));
        }
    }

    if (COUNT && a.counters) {
        unsigned long long ni = n_inter, nv = n_visit;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            ni += __shfl_xor_sync(0xffffffffu, ni, o);
            nv += __shfl_xor_sync(0xffffffffu, nv, o);
        }
        if (lane == 0) {
            atomicAdd(&a.counters[0], ni);
            atomicAdd(&a.counters[1], nv);
            atomicAdd(&a.counters[2], (unsigned long long)n_tests);
        }
    }
    __syncwarp(); // every lane is done with the stack and the child buffers of this walk
#ifndef NB_BH_TRAV_DYNAMIC
    if (a.cta_cost && lane == 0) atomicAdd(&a.cta_cost[cta_id], n_tests);
#endif
    if (valid) {
    // ---- epilogue: the direct-sum kernel's (nb_direct_f64.cu): kick2 / kick1 / drift
    const unsigned idx = a.perm[k];
    const BhParams &p = a.p;
    const double a3[3] = {p.G * ax, p.G * ay, p.G * az};
    if (p.flags & NB_F_STORE_ACC) p.acc[idx] = make_double4(a3[0], a3[1], a3[2], 0.0);
    if (p.flags & (NB_F_KICK2 | NB_F_KICK1_DRIFT)) {
        const double4 V = p.vel[idx];
        double v3[3] = {V.x, V.y, V.z};
        double x3[3] = {me.x, me.y, me.z};
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            if (p.flags & NB_F_KICK2) v3[c] = __dadd_rn(v3[c], __dmul_rn(__dmul_rn(a3[c], p.dt), 0.5));
            if (p.flags & NB_F_KICK1_DRIFT) {
                v3[c] = __dadd_rn(v3[c], __dmul_rn(__dmul_rn(a3[c], p.dt), 0.5));
                x3[c] = __dadd_rn(x3[c], __dmul_rn(v3[c], p.dt));
            }
        }
        p.vel[idx] = make_double4(v3[0], v3[1], v3[2], 0.0);
        if (p.flags & NB_F_KICK1_DRIFT) {
            double *qn = reinterpret_cast<double *>(p.q_next + idx);
            qn[0] = x3[0]; qn[1] = x3[1]; qn[2] = x3[2]; // mass was placed at upload
        }
    }
    } // valid
#ifndef NB_BH_TRAV_DYNAMIC
    break;
#endif
    } // next 32 bodies
}

// Start order of the walk's CTAs: heaviest first, by the records each CTA's bodies tested in the
// previous evaluation (the bodies of CTA i are almost the same from one step to the next).  The
// hardware hands out CTAs in index order; a walk through a dense core takes several times the
// average ~0.8 ms, and when such CTAs come late the other SMs idle until they are done (ncu: SMs
// idle 4 % of the kernel at 1M bodies, 8 % for one rank of eight at 16M).  Measured: 400k Plummer
// 2.08 -> 1.98 ms; from 1M up nothing (see the launch site), so it runs for 151k..600k bodies.  A stable counting sort into 8 cost
// classes (neighbours in a class stay neighbours: they share the tree's lines in L2), one CTA.
// The order in which walks run does not touch any result.
constexpr int ORDER_BINS = 8;
__global__ void __launch_bounds__(1024)
    cta_order_kernel(unsigned *__restrict__ cost, unsigned *__restrict__ order,
                     const unsigned *__restrict__ flags, unsigned n_own_static, int sliced)
{
    __shared__ unsigned s_warp[32];
    __shared__ unsigned s_max, s_base, s_total;
    const unsigned n_own = sliced ? flags[2] : n_own_static;
    const unsigned n = (n_own + TRAV_WARPS * 32 - 1) / (TRAV_WARPS * 32);
    const unsigned t = threadIdx.x, lane = t & 31, w = t >> 5;
    const unsigned per = (n + 1023u) / 1024u;
    const unsigned lo = min(n, t * per), hi = min(n, lo + per);
    unsigned mx = 0;
    for (unsigned i = lo; i < hi; ++i) mx = max(mx, cost[i]);
    mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0) s_warp[w] = mx;
    __syncthreads();
    if (w == 0) {
        const unsigned m = __reduce_max_sync(0xffffffffu, s_warp[lane]);
        if (lane == 0) s_max = m;
    }
    __syncthreads();
    const unsigned long long top = (unsigned long long)s_max + 1ull;
    auto bin_of = [&](unsigned c) { // class 0 = the heaviest
        return (unsigned)(ORDER_BINS - 1) - (unsigned)((unsigned long long)c * ORDER_BINS / top);
    };
    unsigned c[ORDER_BINS];
#pragma unroll
    for (int b = 0; b < ORDER_BINS; ++b) c[b] = 0;
    for (unsigned i = lo; i < hi; ++i) {
        const unsigned b = bin_of(cost[i]);
#pragma unroll
        for (int k = 0; k < ORDER_BINS; ++k) c[k] += (b == (unsigned)k);
    }
    if (t == 0) s_base = 0;
    __syncthreads();
    // exclusive scan, class-major, thread-minor
    unsigned off[ORDER_BINS];
#pragma unroll
    for (int b = 0; b < ORDER_BINS; ++b) {
        unsigned x = c[b];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned y = __shfl_up_sync(0xffffffffu, x, o);
            if ((int)lane >= o) x += y;
        }
        if (lane == 31) s_warp[w] = x;
        __syncthreads();
        if (w == 0) {
            const unsigned v = s_warp[lane];
            unsigned z = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned y = __shfl_up_sync(0xffffffffu, z, o);
                if ((int)lane >= o) z += y;
            }
            s_warp[lane] = z - v;
            if (lane == 31) s_total = z; // the class total
        }
        __syncthreads();
        off[b] = s_base + s_warp[w] + x - c[b];
        __syncthreads();
        if (t == 0) s_base += s_total;
        __syncthreads();
    }
    for (unsigned i = lo; i < hi; ++i) {
        const unsigned b = bin_of(cost[i]);
        unsigned pos = 0;
#pragma unroll
        for (int k = 0; k < ORDER_BINS; ++k)
            if (b == (unsigned)k) pos = off[k]++;
        order[pos] = i;
        cost[i] = 0;
    }
}

__global__ void reset_kernel(unsigned *level_count, unsigned *flags, double *bbox, unsigned *first,
                             unsigned *count, unsigned char *level, unsigned n,
                             unsigned long long *counters, int reset_counters)
{
    if (threadIdx.x < BH_MAX_LEVEL + 2) level_count[threadIdx.x] = threadIdx.x == 0 ? 1u : 0u;
    if (threadIdx.x == 0) {
        bbox[0] = 0.0;
        first[0] = 0;
        count[0] = n;
        level[0] = 0;
        flags[2] = 0;
        flags[4] = 0; // the walk's work counter (traverse_kernel)
        if (reset_counters) counters[0] = counters[1] = counters[2] = 0;
    }
}

} // namespace

// ------------------------------------------------------------------ host ----
unsigned long long bh_node_cap_of(size_t n) { return 2ull * n + 4096ull; }
// nb_init refuses n with bh_node_cap_of(n) > BH_CODE_ID: ids share a word with the child count
static unsigned bh_node_cap(unsigned n) { return (unsigned)bh_node_cap_of(n); }

size_t bh_bytes(unsigned n)
{
    const size_t cap = bh_node_cap(n), nn = n ? n : 1;
    return nn * (2 * 8 + 2 * 4 + 32 + 4) + cap * (32 + 32 + 4 + 4 + 1 + 384 + 4) +
           (size_t)256 * bh_sort_blocks(n) * 4 + 4096;
}

#define BH_TRY(call)                       \
    do {                                   \
        cudaError_t _e = (call);           \
        if (_e != cudaSuccess) return _e;  \
    } while (0)

cudaError_t bh_alloc(BhTree *t, unsigned n, unsigned n_keys, unsigned parts)
{
    const size_t nn = n ? n : 1;
    t->n = n;
    t->n_keys = n_keys > n ? n_keys : n;
    t->node_cap = bh_node_cap(n);
    t->sort_blocks = bh_sort_blocks(n);
    const size_t nk = t->n_keys ? t->n_keys : 1;
    for (int b = 0; b < 2; ++b) {
        BH_TRY(cudaMalloc(&t->key[b], nk * sizeof(uint64_t)));
        BH_TRY(cudaMalloc(&t->val[b], nk * sizeof(unsigned)));
    }
    BH_TRY(cudaMalloc(&t->bbox_parts, (parts ? parts : 1) * sizeof(double)));
    BH_TRY(cudaMemset(t->bbox_parts, 0, (parts ? parts : 1) * sizeof(double)));
    BH_TRY(cudaMalloc(&t->spos, nn * sizeof(double4)));
    BH_TRY(cudaMalloc(&t->own, nn * sizeof(unsigned)));
    BH_TRY(cudaMalloc(&t->own_counts, (nn / OWN_TILE + 2) * sizeof(unsigned)));
    {
        const size_t ctas = nn / (TRAV_WARPS * 32) + 2;
        BH_TRY(cudaMalloc(&t->cta_cost, ctas * sizeof(unsigned)));
        BH_TRY(cudaMalloc(&t->cta_order, ctas * sizeof(unsigned)));
        BH_TRY(cudaMemset(t->cta_cost, 0, ctas * sizeof(unsigned)));
    }
    BH_TRY(cudaMalloc(&t->com, (size_t)t->node_cap * sizeof(double4)));
    BH_TRY(cudaMalloc(&t->child, (size_t)t->node_cap * 8 * sizeof(int)));
    BH_TRY(cudaMalloc(&t->first, (size_t)t->node_cap * sizeof(unsigned)));
    BH_TRY(cudaMalloc(&t->count, (size_t)t->node_cap * sizeof(unsigned)));
    BH_TRY(cudaMalloc(&t->level, (size_t)t->node_cap));
    BH_TRY(cudaMalloc(&t->cblock, (size_t)t->node_cap * 8 * sizeof(ChildRec)));
    BH_TRY(cudaMalloc(&t->pslot, (size_t)t->node_cap * sizeof(unsigned)));
    BH_TRY(cudaMalloc(&t->level_count, (BH_MAX_LEVEL + 2) * sizeof(unsigned)));
    BH_TRY(cudaMalloc(&t->flags, 8 * sizeof(unsigned)));
    BH_TRY(cudaMemset(t->flags, 0, 8 * sizeof(unsigned)));
    BH_TRY(cudaMalloc(&t->bbox, 2 * sizeof(double)));
    BH_TRY(cudaMalloc(&t->hist, ((size_t)256 * t->sort_blocks + 256) * sizeof(unsigned)));
    BH_TRY(cudaMalloc(&t->counters, 4 * sizeof(unsigned long long)));
    BH_TRY(cudaMemset(t->counters, 0, 4 * sizeof(unsigned long long)));
    for (int k = 0; k < BH_STAGES; ++k) {
        BH_TRY(cudaEventCreate(&t->evb[k]));
        BH_TRY(cudaEventCreate(&t->eve[k]));
    }
    return cudaSuccess;
}

void bh_free(BhTree *t)
{
    void *bufs[] = {t->key[0], t->key[1], t->val[0], t->val[1], t->spos, t->own, t->own_counts, t->cta_cost, t->cta_order, t->com, t->child,
                    t->first, t->count, t->level, t->cblock, t->pslot, t->level_count, t->flags, t->bbox, t->hist,
                    t->counters, t->bbox_parts};
    for (void *b : bufs)
        if (b) cudaFree(b);
    for (int k = 0; k < BH_STAGES; ++k) {
        if (t->evb[k]) cudaEventDestroy(t->evb[k]);
        if (t->eve[k]) cudaEventDestroy(t->eve[k]);
    }
    *t = BhTree();
}

cudaError_t bh_morton_order(BhTree *t, const double4 *pos, unsigned n, unsigned *order,
                            cudaStream_t s, unsigned *launches)
{
    if (n == 0) return cudaSuccess;
    unsigned nl = 0;
    const unsigned nblk = (n + BH_THREADS - 1) / BH_THREADS;
    const unsigned wide = nblk < 148u * 8u ? nblk : 148u * 8u;
    reset_kernel<<<1, 32, 0, s>>>(t->level_count, t->flags, t->bbox, t->first, t->count, t->level,
                                  n, t->counters, 0);
    bbox_kernel<<<wide, BH_THREADS, 0, s>>>(pos, n, reinterpret_cast<unsigned long long *>(t->bbox));
    morton_kernel<<<nblk, BH_THREADS, 0, s>>>(pos, n, t->bbox, t->key[0], t->val[0]);
    nl += 3;
    BH_TRY(bh_radix_sort(t, n, 3 * BH_MAX_LEVEL, s, &nl));
    BH_TRY(cudaMemcpyAsync(order, t->val[0], (size_t)n * sizeof(unsigned), cudaMemcpyDeviceToDevice, s));
    if (launches) *launches += nl;
    return cudaGetLastError();
}

// stage boundaries for nb_tree_stats (plain launches only: never inside a graph capture)
enum { ST_KEYS = 0, ST_SORT, ST_MERGE, ST_ORDER, ST_BUILD, ST_COM, ST_TRAV };
static cudaError_t stage_begin(BhTree *t, const BhParams &p, int k, cudaStream_t s)
{
    if (p.time_stages) t->stage_seen[k] = true;
    return p.time_stages ? cudaEventRecord(t->evb[k], s) : cudaSuccess;
}
static cudaError_t stage_end(BhTree *t, const BhParams &p, int k, cudaStream_t s)
{
    return p.time_stages ? cudaEventRecord(t->eve[k], s) : cudaSuccess;
}

// everything after the sorted (key, index) sequence exists in key[t->sorted] / val[t->sorted]:
// bodies in Morton order, own list, build, centre of mass, traversal with the fused epilogue
static cudaError_t bh_after_sort(BhTree *t, const BhParams &p, cudaStream_t s, unsigned *launches)
{
    const unsigned n = p.n;
    unsigned nl = 0;
    const unsigned nblk = (n + BH_THREADS - 1) / BH_THREADS;
    const unsigned wide = nblk < 148u * 8u ? nblk : 148u * 8u; // grid-stride kernels: 8 CTAs per SM
    const uint64_t *skey = t->key[t->sorted];
    const unsigned *perm = t->val[t->sorted];
    BH_TRY(stage_begin(t, p, ST_ORDER, s));
    gather_sorted_kernel<<<nblk, BH_THREADS, 0, s>>>(p.q_cur, perm, n, t->spos);
    ++nl;
    const bool sliced = !(p.own_begin == 0 && p.own_end >= n);
    if (sliced) {
        const unsigned tiles = (n + OWN_TILE - 1) / OWN_TILE;
        own_count_kernel<<<tiles, OWN_TILE, 0, s>>>(perm, n, p.own_begin, p.own_end, t->own_counts);
        BH_TRY(bh_exclusive_scan(t->own_counts, tiles, &t->flags[2], s));
        own_write_kernel<<<tiles, OWN_TILE, 0, s>>>(perm, n, p.own_begin, p.own_end,
                                                    t->own_counts, t->own);
        nl += 3;
    }
    BH_TRY(stage_end(t, p, ST_ORDER, s));
    if (n >= 2) {
        BH_TRY(stage_begin(t, p, ST_BUILD, s));
        BuildArgs ba{skey, t->first, t->count, t->level, t->child, t->level_count, t->flags,
                     t->node_cap, t->spos, t->cblock, t->pslot};
        // a level holds at most n/2 nodes, 8 lanes each
        const unsigned bgrid = min(wide, (n / 2 * 8 + BH_THREADS - 1) / BH_THREADS + 1);
        for (int L = 0; L < BH_MAX_LEVEL; ++L) build_level_kernel<<<bgrid, BH_THREADS, 0, s>>>(ba, L);
        BH_TRY(stage_end(t, p, ST_BUILD, s));
        BH_TRY(stage_begin(t, p, ST_COM, s));
        const unsigned cgrid = min(wide, (n / 2 + BH_THREADS - 1) / BH_THREADS + 1);
        for (int L = BH_MAX_LEVEL; L >= 0; --L)
            com_level_kernel<<<cgrid, BH_THREADS, 0, s>>>(t->spos, t->child, t->first, t->count,
                                                          t->level_count, t->com, L, t->node_cap, t->bbox,
                                                          t->cblock, t->pslot, p.theta, p.eps2);
        BH_TRY(stage_end(t, p, ST_COM, s));
        nl += 2 * BH_MAX_LEVEL + 1;
    }
    static const bool in_order = getenv("NB_BH_BODY_ORDER") != nullptr; // A/B: CTAs in body order
    const unsigned max_own_bodies =
        sliced ? (p.own_end > p.own_begin ? min(p.own_end, n) - min(p.own_begin, n) : 0u) : n;
    // worth it for a few resident grids' worth of bodies (one grid = 151k): 400k Plummer 2.08 ->
    // 1.98 ms.  From 1M up the end of the kernel is a small part of it and the order costs as much
    // as it gains (1M in the bench 4.85 -> 4.91 ms, uniform cube 3.07 -> 3.12, 16M 81.7 -> 82.0,
    // the slowest of eight ranks at 16M unchanged); NB_BH_HEAVY_FIRST=1 forces it on
    static const bool forced = getenv("NB_BH_HEAVY_FIRST") != nullptr;
    const bool lpt = !in_order && !p.count_visits && max_own_bodies > 148u * 8u * TRAV_WARPS * 32u &&
                     (forced || max_own_bodies <= 600000u);
    BH_TRY(stage_begin(t, p, ST_TRAV, s));
    if (lpt) { // more CTAs than fit at once: decide who starts first
        cta_order_kernel<<<1, 1024, 0, s>>>(t->cta_cost, t->cta_order, t->flags, n, sliced ? 1 : 0);
        ++nl;
    }
    TravArgs ta;
    ta.spos = t->spos; ta.com = t->com; ta.cblock = t->cblock;
    ta.first = t->first; ta.count = t->count; ta.perm = perm;
    ta.own = sliced ? t->own : nullptr;
    ta.cta_order = lpt ? t->cta_order : nullptr;
    ta.cta_cost = lpt ? t->cta_cost : nullptr;
    ta.bbox = t->bbox; ta.flags = t->flags; ta.counters = t->counters;
    ta.n_own_static = n;
    ta.p = p;
    const unsigned max_own = sliced ? (p.own_end > p.own_begin ? min(p.own_end, n) - min(p.own_begin, n) : 0u) : n;
    if (max_own) {
        unsigned tgrid = (max_own + TRAV_WARPS * 32 - 1) / (TRAV_WARPS * 32);
#ifdef NB_BH_TRAV_DYNAMIC
        tgrid = min(tgrid, 148u * (unsigned)NB_BH_TRAV_MINB); // resident: every warp loops over the counter
#endif
#ifdef NB_BH_ALWAYS_SELF
        const bool self = true;
#else
        // m / eps2^1.5 must stay finite for the own body's zero term to vanish: any softening a run
        // would use (the reference's is 0.01) is far above the bound
        const bool self = !(p.eps2 >= 1e-60);
#endif
        static const int force_pf = getenv("NB_BH_PREFETCH") ? atoi(getenv("NB_BH_PREFETCH")) : -1;
        const bool pf = force_pf >= 0 ? force_pf != 0 : false;
        if (p.count_visits) traverse_kernel<true, true, false><<<tgrid, TRAV_WARPS * 32, 0, s>>>(ta);
        else if (self) traverse_kernel<false, true, false><<<tgrid, TRAV_WARPS * 32, 0, s>>>(ta);
        else if (pf) traverse_kernel<false, false, true><<<tgrid, TRAV_WARPS * 32, 0, s>>>(ta);
        else traverse_kernel<false, false, false><<<tgrid, TRAV_WARPS * 32, 0, s>>>(ta);
        ++nl;
    }
    BH_TRY(stage_end(t, p, ST_TRAV, s));
    t->stages_timed = p.time_stages != 0;
    if (launches) *launches += nl;
    return cudaGetLastError();
}

cudaError_t bh_evaluate(BhTree *t, const BhParams &p, cudaStream_t s, unsigned *launches)
{
    const unsigned n = p.n;
    unsigned nl = 0;
    if (n == 0) return cudaSuccess;
    const unsigned nblk = (n + BH_THREADS - 1) / BH_THREADS;
    const unsigned wide = nblk < 148u * 8u ? nblk : 148u * 8u;
    for (bool &b : t->stage_seen) b = false;
    t->stages_timed = false;
    BH_TRY(stage_begin(t, p, ST_KEYS, s));
    reset_kernel<<<1, 32, 0, s>>>(t->level_count, t->flags, t->bbox, t->first, t->count, t->level,
                                  n, t->counters, p.count_visits);
    bbox_kernel<<<wide, BH_THREADS, 0, s>>>(p.q_cur, n, reinterpret_cast<unsigned long long *>(t->bbox));
    morton_kernel<<<nblk, BH_THREADS, 0, s>>>(p.q_cur, n, t->bbox, t->key[0], t->val[0]);
    nl += 3;
    BH_TRY(stage_end(t, p, ST_KEYS, s));
    if (p.emulate_runs > 1) {
        // single-rank rehearsal of the multi-rank sort (tests on a one-GPU box): the key sequence
        // is cut into `emulate_runs` slices, each sorted on its own, the runs merged
        const unsigned runs = (unsigned)p.emulate_runs;
        unsigned per = (n + runs - 1) / runs;
        per = (per + 127u) / 128u * 128u;
        if ((unsigned long long)per * runs > t->n_keys) return cudaErrorInvalidValue;
        BH_TRY(stage_begin(t, p, ST_SORT, s));
        for (unsigned r = 0; r < runs; ++r) {
            morton_slice_kernel<<<(per + BH_THREADS - 1) / BH_THREADS, BH_THREADS, 0, s>>>(
                p.q_cur, n, r * per, per, t->bbox, t->key[0], t->val[0]);
            ++nl;
            BH_TRY(bh_radix_sort(t, per, 3 * BH_MAX_LEVEL + 1, s, &nl, r * per));
        }
        BH_TRY(stage_end(t, p, ST_SORT, s));
        BH_TRY(stage_begin(t, p, ST_MERGE, s));
        BH_TRY(bh_merge_runs(t, runs * per, per, s, &nl));
        BH_TRY(stage_end(t, p, ST_MERGE, s));
    } else {
        BH_TRY(stage_begin(t, p, ST_SORT, s));
        BH_TRY(bh_radix_sort(t, n, 3 * BH_MAX_LEVEL, s, &nl));
        BH_TRY(stage_end(t, p, ST_SORT, s));
        t->sorted = 0;
    }
    if (launches) *launches += nl;
    return bh_after_sort(t, p, s, launches);
}

cudaError_t bh_slice_max(BhTree *t, const double4 *q, unsigned n, unsigned rank, unsigned per,
                         cudaStream_t s, unsigned *launches)
{
    const unsigned long long b = (unsigned long long)rank * per;
    const unsigned begin = (unsigned)(b < n ? b : n), count = min(per, n - begin);
    BH_TRY(cudaMemsetAsync(t->bbox_parts + rank, 0, sizeof(double), s));
    if (count) {
        const unsigned nblk = (count + BH_THREADS - 1) / BH_THREADS;
        bbox_kernel<<<nblk < 148u * 8u ? nblk : 148u * 8u, BH_THREADS, 0, s>>>(
            q + begin, count, reinterpret_cast<unsigned long long *>(t->bbox_parts + rank));
        if (launches) ++*launches;
    }
    return cudaGetLastError();
}

cudaError_t bh_slice_sort(BhTree *t, const BhParams &p, unsigned rank, unsigned world, unsigned per,
                          cudaStream_t s, unsigned *launches)
{
    unsigned nl = 0;
    for (bool &b : t->stage_seen) b = false;
    t->stages_timed = false;
    BH_TRY(stage_begin(t, p, ST_KEYS, s));
    reset_kernel<<<1, 32, 0, s>>>(t->level_count, t->flags, t->bbox, t->first, t->count, t->level,
                                  p.n, t->counters, p.count_visits);
    bbox_combine_kernel<<<1, 32, 0, s>>>(t->bbox_parts, world, t->bbox);
    const unsigned begin = rank * per;
    morton_slice_kernel<<<(per + BH_THREADS - 1) / BH_THREADS, BH_THREADS, 0, s>>>(
        p.q_cur, p.n, begin, per, t->bbox, t->key[0], t->val[0]);
    nl += 3;
    BH_TRY(stage_end(t, p, ST_KEYS, s));
    BH_TRY(stage_begin(t, p, ST_SORT, s));
    BH_TRY(bh_radix_sort(t, per, 3 * BH_MAX_LEVEL + 1, s, &nl, begin)); // + 1: the padding key's top bit
    BH_TRY(stage_end(t, p, ST_SORT, s));
    if (launches) *launches += nl;
    return cudaGetLastError();
}

cudaError_t bh_finish(BhTree *t, const BhParams &p, unsigned world, unsigned per, cudaStream_t s,
                      unsigned *launches)
{
    BH_TRY(stage_begin(t, p, ST_MERGE, s));
    BH_TRY(bh_merge_runs(t, world * per, per, s, launches));
    BH_TRY(stage_end(t, p, ST_MERGE, s));
    return bh_after_sort(t, p, s, launches);
}

cudaError_t bh_stats(BhTree *t, BhStats *out, cudaStream_t s)
{
    unsigned lc[BH_MAX_LEVEL + 2], fl[4];
    double bb[2];
    unsigned long long cn[3];
    BH_TRY(cudaMemcpyAsync(lc, t->level_count, sizeof(lc), cudaMemcpyDeviceToHost, s));
    BH_TRY(cudaMemcpyAsync(fl, t->flags, sizeof(fl), cudaMemcpyDeviceToHost, s));
    BH_TRY(cudaMemcpyAsync(bb, t->bbox, sizeof(double), cudaMemcpyDeviceToHost, s));
    BH_TRY(cudaMemcpyAsync(cn, t->counters, sizeof(cn), cudaMemcpyDeviceToHost, s));
    BH_TRY(cudaStreamSynchronize(s));
    out->nodes = 0;
    out->depth = 0;
    for (int l = 0; l <= BH_MAX_LEVEL; ++l) {
        out->nodes += lc[l];
        if (lc[l]) out->depth = (unsigned)l + 1; // the reference counts the root as depth 1
    }
    out->pool_overflow = fl[0];
    out->stack_overflow = fl[1];
    out->root_border = 2.0 * bb[0];
    out->interactions = cn[0];
    out->visits = cn[1];
    out->child_tests = cn[2];
    for (int k = 0; k < BH_STAGES; ++k) {
        float ms = 0.f;
        if (t->stages_timed && t->stage_seen[k]) BH_TRY(cudaEventElapsedTime(&ms, t->evb[k], t->eve[k]));
        out->stage_ms[k] = ms;
    }
    return cudaSuccess;
}

} // namespace nb
