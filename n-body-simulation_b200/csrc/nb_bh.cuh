// nb_bh.cuh -- Barnes-Hut force evaluation for sm_100a: Morton keys, radix
// sort, level-wise octree build, bottom-up centre of mass, warp-cooperative
// THETA traversal with a shared-memory stack (nb_bh.cu, nb_bh_sort.cu).
//
// Stands in for the reference's tree pipeline (serial/barnes-hut.c:114-434,
// cuda/barnes-hut.cu:154-629).  The CELL SET is the reference's: cube
// [-max/2, max/2]^3 centred on the origin with max = 2 max|coord|
// (serial/barnes-hut.c:326-338), octant rule `>=` goes high, index z*4+y*2+x
// (:114-133), a branch exists iff its cell holds >= 2 bodies, a leaf is one
// body; the reference inserts bodies one by one with locks, here the same
// cells come out of a sort.  FP64 throughout, like the reference.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nb {

constexpr int BH_MAX_LEVEL = 21; // 3 x 21 = 63 key bits; deeper cells become buckets

// A child as the traversal sees it: 48 bytes, three 16-byte words, so a node's block of eight is
// 384 bytes = three 128-byte lines (round 1 padded the record to 64 bytes: a quarter of the tree's
// bytes, written by the build and read by every walk that opens the cell, were padding).
struct __align__(16) ChildRec {
    double4 c;  // body: x,y,z,m; cell: centre of mass + mass
    double b2;  // cell: acceptance threshold border^2 / THETA^2 + eps2 (tested against dist^2 + eps2); body: -1
    int ent;    // >= 0 body (sorted index), <= -2 cell -(id + 2)
    int code;   // cell: what the walk pushes to open it -- id | (children - 1) << 27 | bucket << 30
};
static_assert(sizeof(ChildRec) == 48, "child record: three 16-byte words");
constexpr unsigned BH_BLOCK_WORDS = 8 * sizeof(ChildRec) / 16; // 16-byte words per node block (24)
constexpr unsigned BH_CODE_ID = 0x07ffffffu; // 27 bits of node id
constexpr int BH_CODE_NC_SHIFT = 27;         // 3 bits: non-empty children - 1
constexpr unsigned BH_CODE_BUCKET = 1u << 30; // a finest-level cell: bodies summed directly

// Device buffers of one engine's tree.  Bodies are referred to by their rank k
// in Morton order ("sorted index"); perm[k] is the body's index in the state
// arrays.
struct BhTree {
    unsigned n = 0;        // real bodies
    unsigned n_keys = 0;   // key / val capacity: n, or (multi-rank) the padded slices of all ranks
    int sorted = 0;        // which of key[.] / val[.] holds the sorted sequence of the evaluation
    double *bbox_parts = nullptr; // multi-rank: max |coord| of every rank's slice (all-gathered)
    unsigned node_cap = 0; // internal-node pool size
    // per body
    uint64_t *key[2] = {nullptr, nullptr}; // Morton keys, ping-pong for the sort
    unsigned *val[2] = {nullptr, nullptr}; // body index, ping-pong; val[0] ends as perm
    double4 *spos = nullptr;               // x,y,z,m in Morton order
    unsigned *own = nullptr;               // sorted indices this rank integrates, Morton order
    unsigned *own_counts = nullptr;        // per-tile member counts / bases of the compaction
    // the walk's schedule: records tested by each 128-body CTA in the last evaluation, and the
    // order (heaviest first) in which the next one starts them
    unsigned *cta_cost = nullptr, *cta_order = nullptr;
    // per internal node
    double4 *com = nullptr;   // centre of mass x,y,z + total mass
    int *child = nullptr;     // [node][8]: -1 empty, >= 0 body (sorted index), <= -2 node -(c+2)
    unsigned *first = nullptr, *count = nullptr; // the node's bodies: sorted range
    unsigned char *level = nullptr;              // depth (root = 0)
    // what the traversal reads: per node one 384-byte block of up to 8 child records
    // (non-empty children packed to the front, octant order kept), each carrying
    // everything the walk needs about the child -- so opening a cell is ONE round trip
    struct ChildRec *cblock = nullptr; // [node][8]
    unsigned *pslot = nullptr;         // node -> index of its record in the parent's block
    // bookkeeping (device)
    unsigned *level_count = nullptr; // [BH_MAX_LEVEL + 2] nodes per level
    unsigned *flags = nullptr;       // [0] node-pool overflow, [1] stack overflow, [2] own count, [3] root code
    double *bbox = nullptr;          // [0] max |coord| (as ordered bits during the reduction)
    // radix-sort scratch
    unsigned *hist = nullptr; // [256][sort_blocks]
    unsigned sort_blocks = 0;
    // visit counters for the bench (optional): [0] interactions, [1] node visits,
    // [2] records tested by a warp (each costs the walk 17 FP64 warp instructions)
    unsigned long long *counters = nullptr;
    // stage boundaries of the last evaluation issued as plain launches (BhParams::time_stages)
    cudaEvent_t evb[7] = {}, eve[7] = {};
    bool stage_seen[7] = {};
    bool stages_timed = false;
};
constexpr int BH_STAGES = 7; // keys, sort, merge (multi-rank), order, build, centre of mass, traversal

struct BhParams {
    const double4 *q_cur; // positions + mass, state order (all bodies)
    double4 *q_next;      // next positions (bodies of this rank written)
    double4 *vel;         // velocities, state order
    double4 *acc;         // stored accelerations, state order, or nullptr
    unsigned n;           // real bodies
    unsigned own_begin, own_end; // state-order slice integrated by this rank
    unsigned flags;       // NB_F_* of nb_common.cuh
    double dt, eps2, G, theta;
    int count_visits;     // 1: accumulate tree.counters
    int time_stages;      // 1: record the stage events (plain launches only, never while capturing)
    int emulate_runs;     // > 1 (single rank): sort as that many slices + merge, like a multi-rank engine
};

size_t bh_bytes(unsigned n);
unsigned long long bh_node_cap_of(size_t n); // internal-node pool size for n bodies
// n_keys >= n: capacity of the key / index arrays (padded slices of all ranks); parts: ranks
cudaError_t bh_alloc(BhTree *t, unsigned n, unsigned n_keys = 0, unsigned parts = 1);
void bh_free(BhTree *t);
// One force evaluation: bbox -> keys -> sort -> build -> COM -> traversal with the
// kick/drift epilogue selected by p.flags.  Returns the number of kernel launches
// through *launches.
cudaError_t bh_evaluate(BhTree *t, const BhParams &p, cudaStream_t s, unsigned *launches);
// The same evaluation for a multi-rank engine, cut where the ranks exchange data:
//   bh_slice_max   : max |coord| of the own slice of `q` -> bbox_parts[rank]      (then all-gather
//                    of bbox_parts; it runs before the positions are exchanged)
//   bh_slice_sort  : cube from all ranks' maxima, keys of the own slice, sorted in place at
//                    [rank * per, (rank + 1) * per) of key[0] / val[0] (padding keys = all ones)
//                    -- needs only the own positions, so it overlaps the all-gather of the others'
//                    (then all-gather of the sorted runs)
//   bh_finish      : merge of the W runs, then order / build / centre of mass / traversal as in
//                    bh_evaluate.  The merged sequence equals the full sort, so does the tree.
cudaError_t bh_slice_max(BhTree *t, const double4 *q, unsigned n, unsigned rank, unsigned per,
                         cudaStream_t s, unsigned *launches);
cudaError_t bh_slice_sort(BhTree *t, const BhParams &p, unsigned rank, unsigned world, unsigned per,
                          cudaStream_t s, unsigned *launches);
cudaError_t bh_finish(BhTree *t, const BhParams &p, unsigned world, unsigned per, cudaStream_t s,
                      unsigned *launches);
// Morton order of n bodies given as x,y,z,m rows: order[k] = row of the body of rank k
// (the keys and the sort of bh_evaluate, without the tree).  Used once per upload to lay
// the state out along the curve, so that rank slices are compact regions of space.
cudaError_t bh_morton_order(BhTree *t, const double4 *pos, unsigned n, unsigned *order,
                            cudaStream_t s, unsigned *launches);
// tree statistics of the last evaluation (host sync): internal nodes, depth,
// root border, overflow flags
struct BhStats {
    unsigned nodes, depth, pool_overflow, stack_overflow;
    double root_border;
    unsigned long long interactions, visits, child_tests;
    float stage_ms[7]; // device time of the stages of the last plainly launched evaluation, else 0
};
cudaError_t bh_stats(BhTree *t, BhStats *out, cudaStream_t s);

// stable LSD radix sort of (key, val) pairs on 8-bit digits (nb_bh_sort.cu);
// result in key[0], val[0].  bits: number of low key bits that matter.
// offset: sort the n pairs starting at key[.] + offset / val[.] + offset in place (own-slice sort)
cudaError_t bh_radix_sort(BhTree *t, unsigned n, int bits, cudaStream_t s, unsigned *launches,
                          unsigned offset = 0);
// key[0] / val[0] hold total / run sorted runs of length `run` (the last may be short): merges
// them; the result is in key[t->sorted] / val[t->sorted]
cudaError_t bh_merge_runs(BhTree *t, unsigned total, unsigned run, cudaStream_t s, unsigned *launches);
unsigned bh_sort_blocks(unsigned n);
// in-place exclusive prefix sum of `len` counters (one CTA); the grand total goes to *total
cudaError_t bh_exclusive_scan(unsigned *a, unsigned len, unsigned *total, cudaStream_t s);

} // namespace nb
