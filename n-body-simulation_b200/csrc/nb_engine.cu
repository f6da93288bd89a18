// nb_engine.cu -- C ABI of the engine (include/nbody_engine.h): state machine
// of the fused leapfrog, device memory, multi-rank all-gather.
//
// Leapfrog fusion (reference loop: serial/exhaustive.c:206,218-250):
//   reference step k -> k+1 :  v += a(x_k) dt/2 ; x_{k+1} = x_k + v dt ; v += a(x_{k+1}) dt/2
// A launch cannot drift and then read everybody's new positions, so the force
// launch at x_k carries kick2 of step k-1 and kick1+drift of step k:
//   SYNCED_NOACC --[F,K1,D]--> MIDSTEP --[F,K2,K1,D]--> MIDSTEP ... --[F,K2,store a]--> SYNCED_ACC
//   SYNCED_ACC   --[K1,D from stored a]--> MIDSTEP
// n steps + a gather with velocities = n+1 force evaluations, the reference's count.
//
// Multi-rank (world > 1): every rank owns an i-slice and holds all positions.
// After a drift the own slices are re-assembled by ncclAllGather on a side
// stream; the next force evaluation is issued as two launches on the compute
// stream -- first the j-tiles of the rank's own slice (already valid), then,
// behind an event wait on the all-gather, everybody else's j-tiles + epilogue
// -- so the exchange over NVLink is hidden behind compute.
#include <dlfcn.h>
#include <new>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <mutex>
#include <vector>

#include "../../include/nbody_engine.h"
#include "nb_aux.cuh"
#include "nb_bh.cuh"
#include "nb_direct.cuh"

using namespace nb;

namespace {

thread_local char g_err[512] = "";

void set_err(const char *what, const char *detail)
{
    snprintf(g_err, sizeof(g_err), "%s: %s", what, detail ? detail : "");
}

#define NB_CUDA(call)                                                  \
    do {                                                               \
        cudaError_t _e = (call);                                       \
        if (_e != cudaSuccess) {                                       \
            set_err(#call, cudaGetErrorString(_e));                    \
            return _e == cudaErrorMemoryAllocation ? NB_E_NOMEM : NB_E_CUDA; \
        }                                                              \
    } while (0)

// ---- NCCL, bound at run time (libnccl.so.2: torch's copy if already loaded,
// else the system one).  Single-rank engines never touch it.
struct nccl_id_t {
    char internal[NB_COMM_ID_BYTES];
};
struct NcclApi {
    void *lib = nullptr;
    int (*GetUniqueId)(nccl_id_t *) = nullptr;
    int (*CommInitRank)(void **, int, nccl_id_t, int) = nullptr;
    int (*AllGather)(const void *, void *, size_t, int, void *, cudaStream_t) = nullptr;
    int (*CommDestroy)(void *) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mutex; // engines of one process start on different host threads

int load_nccl()
{
    std::lock_guard<std::mutex> guard(g_nccl_mutex);
    if (g_nccl.lib) return NB_OK;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *nm : names)
        if ((h = dlopen(nm, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!h) {
        set_err("dlopen(libnccl.so.2)", dlerror());
        return NB_E_COMM;
    }
    NcclApi a;
    a.lib = h;
    a.GetUniqueId = (int (*)(nccl_id_t *))dlsym(h, "ncclGetUniqueId");
    a.CommInitRank = (int (*)(void **, int, nccl_id_t, int))dlsym(h, "ncclCommInitRank");
    a.AllGather = (int (*)(const void *, void *, size_t, int, void *, cudaStream_t))dlsym(
        h, "ncclAllGather");
    a.CommDestroy = (int (*)(void *))dlsym(h, "ncclCommDestroy");
    a.GetErrorString = (const char *(*)(int))dlsym(h, "ncclGetErrorString");
    if (!a.GetUniqueId || !a.CommInitRank || !a.AllGather || !a.CommDestroy) {
        set_err("dlsym(nccl)", "missing symbol");
        return NB_E_COMM;
    }
    g_nccl = a;
    return NB_OK;
}

#define NB_NCCL(call)                                                               \
    do {                                                                            \
        int _r = (call);                                                            \
        if (_r != 0) {                                                              \
            set_err(#call, g_nccl.GetErrorString ? g_nccl.GetErrorString(_r) : "nccl error"); \
            return NB_E_COMM;                                                       \
        }                                                                           \
    } while (0)

enum State { SYNCED_NOACC = 0, SYNCED_ACC = 1, MIDSTEP = 2 };

} // namespace

struct nb_engine {
    nb_config cfg;
    size_t n = 0;          // real bodies
    unsigned unit_bodies = 1; // 4 (FP32 quads) or 1 (FP64)
    size_t unit_bytes = 0;    // bytes of one position unit
    size_t vel_unit_bytes = 0;
    unsigned n_units = 0;     // padded j extent in units
    unsigned slice_units = 0; // units per rank
    unsigned i_begin = 0, i_end = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;      // compute
    cudaStream_t comm_stream = nullptr; // NCCL collectives (world > 1)
    cudaEvent_t ev_compute = nullptr, ev_gather = nullptr; // hand-offs between the two
    bool gather_pending = false; // an all-gather is in flight on comm_stream
    // Barnes-Hut, world > 1: the sort is cut along the rank slices (bh_slice_max / bh_slice_sort /
    // bh_finish); after a drift the ranks' cube maxima and the positions are in flight
    bool bh_pipe = false, bh_pipe_pending = false;
    cudaEvent_t ev_max = nullptr, ev_runs = nullptr;
    int bh_emulate_runs = 0;
    bool step_ended_in_gather = false;
    cudaEvent_t ev_step0 = nullptr, ev_step1 = nullptr, ev_f0 = nullptr, ev_f1 = nullptr;
    cudaEvent_t ev_x0 = nullptr; // side stream, just before the position exchange (ev_gather ends it)
    bool have_x_ev = false;
    bool have_step_ev = false, have_force_ev = false;
    bool time_force = true; // nb_step times only the last evaluation of a multi-step call
    void *q[2] = {nullptr, nullptr};
    int cur = 0;
    void *vel = nullptr, *acc = nullptr;
    double *workspace = nullptr, *partial = nullptr;
    unsigned *counters = nullptr;
    double *stage_pos = nullptr, *stage_vel = nullptr; // host-packing staging, 4 / 3 doubles per row
    // host I/O slices (nb_upload_slice / nb_gather_slice): rank r moves rows [r*io_per, (r+1)*io_per)
    // of the caller's order over PCIe, the rest travels between the GPUs (all-gather of the staging rows)
    size_t io_per = 0;
    unsigned *mass_cls = nullptr; // device: [0] all masses equal (direct sum), [1] bits of (float)m_0, [2..3] of m_0
    double *energy_part = nullptr;
    // frame pipeline (nb_frame_begin / nb_frame_wait), created on first use
    cudaStream_t copy_stream = nullptr;
    double *frame_stage[2] = {nullptr, nullptr};                  // unpacked snapshots, 4n doubles
    cudaEvent_t ev_frame_ready[2] = {nullptr, nullptr};           // snapshot written (compute stream)
    cudaEvent_t ev_frame_done[2] = {nullptr, nullptr};            // copied to the host (copy stream)
    const double *frame_host[2] = {nullptr, nullptr};             // where each slot is headed
    unsigned frame_seq = 0;
    DirectPlan plan;
    BhTree tree;            // NB_ALGO_BARNES_HUT only
    // Barnes-Hut: the device keeps the bodies in the Morton order of the uploaded positions
    // (rank slices = compact regions, warps of neighbours); user_of[k] = caller's index of
    // device body k, nullptr = caller's order
    unsigned *user_of = nullptr;
    double *order_tmp = nullptr; // 4n doubles of scratch for the reordering
    unsigned *morton_tmp = nullptr, *chunk_base = nullptr; // world > 1: dealing the curve out to the ranks
    // NB_DEBUG_OWN_SLICE="W:r" on a one-rank Barnes-Hut engine: lay the state out as W ranks would
    // and walk only rank r's bodies -- what one rank of a W-GPU run does, measured on one GPU
    unsigned dbg_world = 0, dbg_rank = 0, dbg_len = 0;
    bool deal_ready = false;      // chunk_base is on the device (it depends on n, world and the piece count only)
    int relayout_every = 0;       // Barnes-Hut, world > 1: steps between two re-layouts of the state (0: never)
    int steps_since_layout = 0;
    // A Barnes-Hut evaluation is ~75 short launches with fixed arguments per (ping-pong side,
    // epilogue flags): captured once as a CUDA graph, replayed afterwards.
    struct BhGraph {
        int cur;
        unsigned flags;
        bool pipe;
        cudaGraphExec_t exec;
        unsigned launches;
    };
    std::vector<BhGraph> bh_graphs;
    bool bh_graphs_off = false;
    bool bh = false;
    bool count_visits = false;
    bool uniform = false;   // direct sum: all masses equal -> mass-free inner loop
    double uniform_m = 0.0;
    State state = SYNCED_NOACC;
    void *comm = nullptr;
    bool vel_complete = true; // world > 1: do all ranks hold all velocities?
    nb_stats st;
};

namespace {

int set_device(nb_engine *e)
{
    NB_CUDA(cudaSetDevice(e->cfg.device));
    return NB_OK;
}

// The compute stream may not touch gathered data before the collective landed.
int wait_gather(nb_engine *e)
{
    if (!e->gather_pending) return NB_OK;
    NB_CUDA(cudaStreamWaitEvent(e->stream, e->ev_gather, 0));
    e->gather_pending = false;
    return NB_OK;
}

// In-place all-gather of every rank's slice of `buf`, on the side stream, after
// everything enqueued on the compute stream so far.  Returns without waiting:
// consumers call wait_gather().
int all_gather_bytes(nb_engine *e, void *buf, size_t bytes);
int all_gather_units(nb_engine *e, void *buf, size_t unit_bytes)
{
    return all_gather_bytes(e, buf, (size_t)e->slice_units * unit_bytes);
}
// in-place all-gather on the side stream behind everything enqueued on the compute stream so
// far; `done` (optional) is recorded behind it.  Does not touch gather_pending / ev_gather.
int comm_all_gather(nb_engine *e, void *buf, size_t bytes, cudaEvent_t done)
{
    if (e->cfg.world <= 1) return NB_OK;
    const char *send = (const char *)buf + (size_t)e->cfg.rank * bytes;
    NB_CUDA(cudaEventRecord(e->ev_compute, e->stream));
    NB_CUDA(cudaStreamWaitEvent(e->comm_stream, e->ev_compute, 0));
    NB_NCCL(g_nccl.AllGather(send, buf, bytes, /*ncclInt8*/ 0, e->comm, e->comm_stream));
    if (done) NB_CUDA(cudaEventRecord(done, e->comm_stream));
    e->st.collectives++;
    return NB_OK;
}

// rank r contributes bytes [r*bytes, (r+1)*bytes) of buf
int all_gather_bytes(nb_engine *e, void *buf, size_t bytes)
{
    if (e->cfg.world <= 1) return NB_OK;
    const char *send = (const char *)buf + (size_t)e->cfg.rank * bytes;
    NB_CUDA(cudaEventRecord(e->ev_compute, e->stream));
    NB_CUDA(cudaStreamWaitEvent(e->comm_stream, e->ev_compute, 0));
    NB_CUDA(cudaEventRecord(e->ev_x0, e->comm_stream));
    NB_NCCL(g_nccl.AllGather(send, buf, bytes, /*ncclInt8*/ 0, e->comm, e->comm_stream));
    NB_CUDA(cudaEventRecord(e->ev_gather, e->comm_stream));
    e->have_x_ev = true;
    e->gather_pending = true;
    e->step_ended_in_gather = false; // nb_step sets it again for the gather that ends a step
    e->st.collectives++;
    return NB_OK;
}

// After a drift wrote the own slice of q[cur ^ 1]: re-assemble everybody's positions there and
// swap the buffers.  Barnes-Hut (pipe): the cube needs the maximum |coord| over ALL bodies before
// any key can be computed, so every rank reduces its own slice first and those W doubles travel
// ahead of the positions; the keys and the sort of the own slice then run under the big exchange.
int exchange_positions(nb_engine *e)
{
    int rc;
    void *qn = e->q[e->cur ^ 1];
    if (e->bh_pipe) {
        unsigned nl = 0;
        NB_CUDA(bh_slice_max(&e->tree, static_cast<const double4 *>(qn), (unsigned)e->n,
                             (unsigned)e->cfg.rank, e->slice_units, e->stream, &nl));
        e->st.kernel_launches += nl;
        if ((rc = comm_all_gather(e, e->tree.bbox_parts, sizeof(double), e->ev_max))) return rc;
        e->bh_pipe_pending = true;
    }
    if ((rc = all_gather_units(e, qn, e->unit_bytes))) return rc;
    e->cur ^= 1;
    return NB_OK;
}

int launch_kernel(nb_engine *e, int geom, const StepParams &p)
{
    if (e->cfg.precision == NB_PREC_FP32)
        NB_CUDA(launch_direct_f32(e->plan, geom, p, e->cfg.tune_flags, e->stream));
    else
        NB_CUDA(launch_direct_f64(e->plan, geom, p, e->cfg.tune_flags, e->stream));
    e->st.kernel_launches++;
    return NB_OK;
}

// One force launch with the given epilogue flags.  Positions move iff
// NB_F_KICK1_DRIFT: then the own slice of q[next] is written, all-gathered
// and the buffers swap.
int launch_force(nb_engine *e, unsigned flags)
{
    StepParams p;
    memset(&p, 0, sizeof(p));
    p.q_cur = e->q[e->cur];
    p.q_next = e->q[e->cur ^ 1];
    p.vel = e->vel;
    p.acc = e->acc;
    p.workspace = e->workspace;
    p.partial = e->partial;
    p.counters = e->counters;
    p.n_real = (unsigned)e->n;
    p.n_units = e->n_units;
    p.i_begin = e->i_begin;
    p.i_end = e->i_end;
    p.flags = flags;
    p.dt = e->cfg.dt;
    p.eps2 = e->cfg.eps2;
    p.G = e->uniform ? e->cfg.G * e->uniform_m : e->cfg.G;
    p.uniform_mass = e->uniform ? 1u : 0u;
    if (e->time_force) NB_CUDA(cudaEventRecord(e->ev_f0, e->stream));
    int rc;
    if (e->bh) {
        const bool pipe = e->bh_pipe_pending;
        // the tree is built from everybody's positions; what a rank can do before they land is
        // the keys and the sort of its own slice (pipe), once every rank's cube maximum is in
        if (!pipe && (rc = wait_gather(e))) return rc;
        BhParams bp;
        bp.q_cur = static_cast<const double4 *>(p.q_cur);
        bp.q_next = static_cast<double4 *>(p.q_next);
        bp.vel = static_cast<double4 *>(p.vel);
        bp.acc = static_cast<double4 *>(p.acc);
        bp.n = (unsigned)e->n;
        bp.own_begin = e->i_begin;
        bp.own_end = e->i_end;
        if (e->dbg_world) {
            bp.own_begin = e->dbg_rank * e->dbg_len;
            bp.own_end = std::min<unsigned>((e->dbg_rank + 1) * e->dbg_len, (unsigned)e->n);
        }
        bp.flags = flags;
        bp.dt = p.dt; bp.eps2 = p.eps2; bp.G = p.G; bp.theta = e->cfg.theta;
        bp.count_visits = e->count_visits ? 1 : 0;
        bp.time_stages = e->bh_graphs_off ? 1 : 0;
        bp.emulate_runs = e->bh_emulate_runs;
        unsigned nl = 0;
        bool done = false;
        const unsigned per = e->slice_units, wr = (unsigned)e->cfg.world, rk = (unsigned)e->cfg.rank;
        if (pipe) {
            NB_CUDA(cudaStreamWaitEvent(e->stream, e->ev_max, 0));
            NB_CUDA(bh_slice_sort(&e->tree, bp, rk, wr, per, e->stream, &nl));
            // the sorted runs follow the positions on the side stream
            if ((rc = comm_all_gather(e, e->tree.key[0], (size_t)per * sizeof(uint64_t), nullptr))) return rc;
            if ((rc = comm_all_gather(e, e->tree.val[0], (size_t)per * sizeof(unsigned), e->ev_runs))) return rc;
            NB_CUDA(cudaStreamWaitEvent(e->stream, e->ev_runs, 0));
            if ((rc = wait_gather(e))) return rc; // the positions went first on that stream
            e->bh_pipe_pending = false;
        }
        auto evaluate = [&](unsigned *count) -> cudaError_t {
            return pipe ? bh_finish(&e->tree, bp, wr, per, e->stream, count)
                        : bh_evaluate(&e->tree, bp, e->stream, count);
        };
        if (!e->bh_graphs_off) {
            for (const nb_engine::BhGraph &g : e->bh_graphs)
                if (g.cur == e->cur && g.flags == flags && g.pipe == pipe) {
                    NB_CUDA(cudaGraphLaunch(g.exec, e->stream));
                    nl += g.launches;
                    done = true;
                    break;
                }
            if (!done) {
                // thread-local capture: other engines of this process (one host thread each)
                // keep allocating and launching meanwhile
                cudaGraph_t graph = nullptr;
                cudaGraphExec_t exec = nullptr;
                cudaError_t ce = cudaStreamBeginCapture(e->stream, cudaStreamCaptureModeThreadLocal);
                unsigned gl = 0;
                if (ce == cudaSuccess) {
                    ce = evaluate(&gl);
                    const cudaError_t ee = cudaStreamEndCapture(e->stream, &graph);
                    if (ce == cudaSuccess) ce = ee;
                }
                if (ce == cudaSuccess) ce = cudaGraphInstantiate(&exec, graph, 0);
                if (graph) cudaGraphDestroy(graph);
                if (ce == cudaSuccess) {
                    e->bh_graphs.push_back({e->cur, flags, pipe, exec, gl});
                    NB_CUDA(cudaGraphLaunch(exec, e->stream));
                    nl += gl;
                    done = true;
                } else { // no graphs on this engine: fall back to plain launches, loudly in the stats
                    cudaGetLastError();
                    e->bh_graphs_off = true;
                }
            }
        }
        if (!done) NB_CUDA(evaluate(&nl));
        e->st.kernel_launches += nl;
    } else if (e->gather_pending && e->plan.can_split && !(e->cfg.tune_flags & NB_TUNE_NO_OVERLAP)) {
        // positions of the other ranks are still in flight: start on the own slice
        StepParams a = p;
        a.flags = NB_F_PARTIAL_OUT;
        if ((rc = launch_kernel(e, NB_GEOM_OWN, a))) return rc;
        if ((rc = wait_gather(e))) return rc;
        p.flags |= NB_F_PARTIAL_IN;
        if ((rc = launch_kernel(e, NB_GEOM_REST, p))) return rc;
        e->st.split_evaluations++;
    } else {
        if ((rc = wait_gather(e))) return rc;
        if ((rc = launch_kernel(e, NB_GEOM_FULL, p))) return rc;
    }
    if (e->time_force) {
        NB_CUDA(cudaEventRecord(e->ev_f1, e->stream));
        e->have_force_ev = true;
    }
    e->st.force_launches++;
    if (!e->bh) // direct sum: every ordered pair; Barnes-Hut counts on the device (nb_tree_stats)
        e->st.interactions += (unsigned long long)(e->i_end - e->i_begin) * e->unit_bodies *
                              (unsigned long long)e->n_units * e->unit_bodies;
    if (flags & (NB_F_KICK2 | NB_F_KICK1_DRIFT)) e->vel_complete = (e->cfg.world <= 1);
    if (flags & NB_F_KICK1_DRIFT) {
        int rc = exchange_positions(e);
        if (rc) return rc;
    }
    return NB_OK;
}

int close_step(nb_engine *e)
{
    if (e->state != MIDSTEP) return NB_OK;
    int rc = launch_force(e, NB_F_KICK2 | NB_F_STORE_ACC);
    if (rc) return rc;
    e->state = SYNCED_ACC;
    return NB_OK;
}

int ensure_vel_complete(nb_engine *e)
{
    if (e->vel_complete) return NB_OK;
    int rc = all_gather_units(e, e->vel, e->vel_unit_bytes);
    if (rc) return rc;
    e->vel_complete = true;
    return wait_gather(e);
}

// World > 1, Barnes-Hut: the ranks' slices are contiguous in the device order, and equal in
// bodies -- but a contiguous piece of the Morton curve is a region of space, and regions differ in
// work per body.  So the curve is dealt out in pieces, round robin: every rank gets every
// world-th piece and so a sample of every region.  Pieces are runs of 128-body chunks (the slice
// alignment); kDealPieces of them per rank: long enough that a rank's walks stay in a few
// subtrees (128-body pieces cost 13 % on 8 GPUs -- every rank then drags the whole deep tree
// through its L2), short enough to even out a lopsided system.
// user_of[device index] = morton_tmp[rank on the curve].
constexpr unsigned kDealChunk = 128; // = the slice alignment of FP64 engines
constexpr unsigned kDealPieces = 32; // pieces per rank (tune_flags bits 16..23: log2 + 1 overrides)

int deal_curve(nb_engine *e)
{
    const unsigned n = (unsigned)e->n;
    const unsigned world = e->dbg_world ? e->dbg_world : (unsigned)e->cfg.world;
    const unsigned len = e->dbg_world ? e->dbg_len : e->slice_units;
    if (e->deal_ready) {
        NB_CUDA(remap_chunks(e->morton_tmp, e->user_of, n, kDealChunk, e->chunk_base, e->stream));
        return NB_OK;
    }
    const unsigned full = n / kDealChunk, rag = n % kDealChunk;
    std::vector<unsigned> base(full + 1, 0u), fill(world, 0u), cap(world, 0u);
    for (unsigned r = 0; r < world; ++r) {
        const size_t lo = (size_t)r * len;
        cap[r] = lo >= n ? 0u : (unsigned)std::min<size_t>(len, n - lo);
    }
    const unsigned lg = ((unsigned)e->cfg.tune_flags >> 16) & 0xffu;
    const unsigned pieces = lg ? 1u << (lg - 1) : kDealPieces;
    const unsigned run = std::max(1u, len / kDealChunk / pieces); // chunks per piece
    unsigned r = 0, in_run = 0;
    for (unsigned c = 0; c < full; ++c) {
        if (in_run == run) { // piece complete: next rank
            r = (r + 1) % world;
            in_run = 0;
        }
        unsigned tries = 0;
        while (fill[r] + kDealChunk > cap[r] && tries++ < world) { // this slice is full
            r = (r + 1) % world;
            in_run = 0;
        }
        if (fill[r] + kDealChunk > cap[r]) {
            set_err("nb_upload", "internal: slice capacities do not add up");
            return NB_E_INVALID;
        }
        base[c] = r * len + fill[r];
        fill[r] += kDealChunk;
        ++in_run;
    }
    if (rag) { // the last, short chunk closes the one slice whose length is not a multiple of 128
        const unsigned rr = (n - 1) / len;
        base[full] = rr * len + fill[rr];
    }
    NB_CUDA(cudaMemcpyAsync(e->chunk_base, base.data(), base.size() * sizeof(unsigned),
                            cudaMemcpyHostToDevice, e->stream));
    NB_CUDA(remap_chunks(e->morton_tmp, e->user_of, n, kDealChunk, e->chunk_base, e->stream));
    NB_CUDA(cudaStreamSynchronize(e->stream)); // `base` leaves scope
    e->deal_ready = true;
    return NB_OK;
}

// rows of the caller's order this rank moves over PCIe in the *_slice calls
void io_rows(const nb_engine *e, size_t *begin, size_t *count)
{
    const size_t b = std::min(e->n, e->io_per * (size_t)e->cfg.rank);
    *begin = b;
    *count = std::min(e->n - b, e->io_per);
}

int layout_from_stage(nb_engine *e, bool have_vel, bool fresh);

// sliced: pos / vel hold only this rank's io rows; the other ranks' rows arrive over NVLink
int upload(nb_engine *e, const double *pos, const double *vel, bool sliced)
{
    const size_t n = e->n;
    int wrc = wait_gather(e); // an all-gather may still be writing the position buffers
    if (wrc) return wrc;
    if (n && !(sliced && e->cfg.world > 1)) {
        NB_CUDA(cudaMemcpyAsync(e->stage_pos, pos, n * 4 * sizeof(double), cudaMemcpyHostToDevice,
                                e->stream));
        if (vel)
            NB_CUDA(cudaMemcpyAsync(e->stage_vel, vel, n * 3 * sizeof(double),
                                    cudaMemcpyHostToDevice, e->stream));
    } else if (n) {
        size_t b, c;
        io_rows(e, &b, &c);
        if (c) {
            NB_CUDA(cudaMemcpyAsync(e->stage_pos + 4 * b, pos, c * 4 * sizeof(double),
                                    cudaMemcpyHostToDevice, e->stream));
            if (vel)
                NB_CUDA(cudaMemcpyAsync(e->stage_vel + 3 * b, vel, c * 3 * sizeof(double),
                                        cudaMemcpyHostToDevice, e->stream));
        }
        int rc = all_gather_bytes(e, e->stage_pos, e->io_per * 4 * sizeof(double));
        if (!rc && vel) rc = all_gather_bytes(e, e->stage_vel, e->io_per * 3 * sizeof(double));
        if (!rc) rc = wait_gather(e);
        if (rc) return rc;
    }
    return layout_from_stage(e, vel != nullptr, true);
}

// The staged rows (host packing, caller's order, on the device) become the device layout.
// fresh: a new state from the host (classify the masses, clear the tree's overflow flags, reset
// the step counter); otherwise the running state re-laid out (relayout below): nothing but the
// order of the bodies on the device changes, and the host is not waited for.
int layout_from_stage(nb_engine *e, bool have_vel, bool fresh)
{
    const size_t n = e->n;
    const double *vel = have_vel ? e->stage_vel : nullptr;
    // equal masses (the usual N-body initial condition): the FP32 kernel drops m_j from its inner
    // loop and the epilogue multiplies by G*m.  Padding then has to be inert without a zero
    // mass: it is parked so far away that r^-3 underflows to exactly 0.  The masses are
    // classified on the device (mass_class), from the staged rows; the host reads two words.
    const bool f64 = e->cfg.precision != NB_PREC_FP32;
    const bool classify = !e->bh && n > 0 && !(e->cfg.tune_flags & NB_TUNE_NO_UNIFORM) &&
                          (f64 || !(e->cfg.tune_flags & NB_TUNE_SCALAR_FP32));
    if (fresh) {
        NB_CUDA(mass_class(e->stage_pos, (unsigned)n, e->mass_cls, classify, f64, e->stream));
        if (classify) e->st.kernel_launches++;
        // overflow flags of the tree stay set between evaluations (a truncated tree must be caught
        // at the next gather) but not across a new state
        if (e->bh) NB_CUDA(cudaMemsetAsync(e->tree.flags, 0, 2 * sizeof(unsigned), e->stream));
    }
    if (e->user_of) {
        // every rank sees all n bodies, so every rank derives the same order
        unsigned nl = 0;
        NB_CUDA(bh_morton_order(&e->tree, reinterpret_cast<const double4 *>(e->stage_pos), (unsigned)n,
                                e->morton_tmp ? e->morton_tmp : e->user_of, e->stream, &nl));
        if (e->morton_tmp) {
            int rc = deal_curve(e);
            if (rc) return rc;
            ++nl;
        }
        NB_CUDA(permute_rows(e->stage_pos, e->order_tmp, e->user_of, (unsigned)n, 4, e->stream));
        NB_CUDA(cudaMemcpyAsync(e->stage_pos, e->order_tmp, n * 4 * sizeof(double),
                                cudaMemcpyDeviceToDevice, e->stream));
        if (vel) {
            NB_CUDA(permute_rows(e->stage_vel, e->order_tmp, e->user_of, (unsigned)n, 3, e->stream));
            NB_CUDA(cudaMemcpyAsync(e->stage_vel, e->order_tmp, n * 3 * sizeof(double),
                                    cudaMemcpyDeviceToDevice, e->stream));
        }
        e->st.kernel_launches += nl + (vel ? 2 : 1);
    }
    NB_CUDA(pack_bodies(e->cfg.precision, e->stage_pos, vel ? e->stage_vel : nullptr, (unsigned)n,
                        e->n_units, e->q[0], e->q[1], e->vel, e->mass_cls, e->stream));
    e->st.kernel_launches++;
    e->cur = 0;
    e->vel_complete = true;
    e->bh_pipe_pending = false; // all positions are local again: the next evaluation sorts them itself
    e->steps_since_layout = 0;
    if (!fresh) return NB_OK;
    unsigned cls[4] = {0u, 0u, 0u, 0u};
    NB_CUDA(cudaMemcpyAsync(cls, e->mass_cls, sizeof(cls), cudaMemcpyDeviceToHost, e->stream));
    NB_CUDA(cudaStreamSynchronize(e->stream)); // the caller's arrays are free again; cls is in
    float m0;
    double d0;
    memcpy(&m0, &cls[1], sizeof(m0));
    memcpy(&d0, &cls[2], sizeof(d0));
    e->uniform = cls[0] != 0u;
    e->uniform_m = f64 ? d0 : (double)m0;
    e->state = SYNCED_NOACC;
    e->st.steps_done = 0;
    return NB_OK;
}

// Barnes-Hut on several ranks: a rank integrates the bodies it was dealt when the state was laid
// out (pieces of the Morton curve of THAT moment).  As the bodies move, a rank's bodies spread over
// the curve; 32 consecutive ones of its list are then no longer neighbours and their common walk
// tests more records (16M bodies on 8 GPUs: +8 % after 80 steps of a Plummer sphere).  Every
// relayout_every steps the state is therefore laid out again, on the device: velocities
// all-gathered, state unpacked to the staged rows (caller's order) and packed from there exactly
// as nb_upload does -- no value changes, only where it lives and who integrates it.  Collective;
// called between two steps (state MIDSTEP: positions drifted, velocities half-kicked).
int relayout(nb_engine *e)
{
    int rc = wait_gather(e);
    if (rc) return rc;
    if ((rc = ensure_vel_complete(e))) return rc;
    const unsigned n = (unsigned)e->n;
    NB_CUDA(unpack_bodies(e->cfg.precision, e->vel, n, 3, e->stage_vel, e->user_of, e->stream));
    NB_CUDA(unpack_bodies(e->cfg.precision, e->q[e->cur], n, 4, e->stage_pos, e->user_of, e->stream));
    e->st.kernel_launches += 2;
    return layout_from_stage(e, true, false);
}

} // namespace

// =============================================================== C ABI =====
extern "C" {

void nb_config_default(nb_config *cfg)
{
    memset(cfg, 0, sizeof(*cfg));
    cfg->precision = NB_PREC_FP64;
    cfg->algorithm = NB_ALGO_DIRECT;
    cfg->dt = (double)0.01f;
    cfg->eps2 = 0.01;
    cfg->G = 1.0;
    cfg->theta = 0.5;
    cfg->device = 0;
    cfg->rank = 0;
    cfg->world = 1;
}

int nb_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

const char *nb_version(void) { return "nbody_b200 0.1 (sm_100a)"; }

const char *nb_last_error(void) { return g_err; }

const char *nb_strerror(int code)
{
    switch (code) {
    case NB_OK: return "ok";
    case NB_E_INVALID: return "invalid argument";
    case NB_E_CUDA: return "CUDA error";
    case NB_E_NOMEM: return "out of memory";
    case NB_E_COMM: return "communicator (NCCL) error";
    case NB_E_NODEVICE: return "no CUDA device (this engine has no CPU fallback)";
    case NB_E_STATE: return "call not valid in this state";
    default: return "unknown error";
    }
}

int nb_comm_id(void *id_out, size_t bytes)
{
    if (!id_out || bytes < NB_COMM_ID_BYTES) return NB_E_INVALID;
    int rc = load_nccl();
    if (rc) return rc;
    nccl_id_t id;
    NB_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(id_out, &id, NB_COMM_ID_BYTES);
    return NB_OK;
}

void nb_free(nb_engine *e)
{
    if (!e) return;
    cudaSetDevice(e->cfg.device);
    if (e->comm_stream) cudaStreamSynchronize(e->comm_stream);
    if (e->stream) cudaStreamSynchronize(e->stream);
    if (e->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(e->comm);
    bh_free(&e->tree);
    for (const nb_engine::BhGraph &g : e->bh_graphs) cudaGraphExecDestroy(g.exec);
    void *bufs[] = {e->q[0], e->q[1], e->vel, e->acc, e->workspace, e->partial, e->counters,
                    e->stage_pos, e->stage_vel, e->mass_cls, e->energy_part, e->frame_stage[0], e->frame_stage[1],
                    e->user_of, e->order_tmp, e->morton_tmp, e->chunk_base};
    if (e->copy_stream) cudaStreamSynchronize(e->copy_stream);
    for (void *b : bufs)
        if (b) cudaFree(b);
    cudaEvent_t evs[] = {e->ev_x0, e->ev_step0, e->ev_step1, e->ev_f0, e->ev_f1, e->ev_compute, e->ev_gather,
                         e->ev_max, e->ev_runs,
                         e->ev_frame_ready[0], e->ev_frame_ready[1], e->ev_frame_done[0],
                         e->ev_frame_done[1]};
    for (cudaEvent_t ev : evs)
        if (ev) cudaEventDestroy(ev);
    if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    if (e->comm_stream) cudaStreamDestroy(e->comm_stream);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

static int init_impl(nb_engine *e, const double *pos, const double *vel)
{
    const nb_config &c = e->cfg;
    int rc = set_device(e);
    if (rc) return rc;
    cudaDeviceProp prop;
    NB_CUDA(cudaGetDeviceProperties(&prop, c.device));
    if (prop.major < 10) {
        set_err("nb_init", "device is not sm_100 class; this library carries sm_100a code only");
        return NB_E_NODEVICE;
    }
    e->sm_count = prop.multiProcessorCount;
    NB_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    NB_CUDA(cudaEventCreate(&e->ev_step0));
    NB_CUDA(cudaEventCreate(&e->ev_step1));
    NB_CUDA(cudaEventCreate(&e->ev_f0));
    NB_CUDA(cudaEventCreate(&e->ev_f1));
    NB_CUDA(cudaEventCreate(&e->ev_x0));
    if (c.world > 1) {
        NB_CUDA(cudaStreamCreateWithFlags(&e->comm_stream, cudaStreamNonBlocking));
        NB_CUDA(cudaEventCreateWithFlags(&e->ev_compute, cudaEventDisableTiming));
        NB_CUDA(cudaEventCreate(&e->ev_gather)); // timed: a step ends when its all-gather landed
        NB_CUDA(cudaEventCreateWithFlags(&e->ev_max, cudaEventDisableTiming));
        NB_CUDA(cudaEventCreateWithFlags(&e->ev_runs, cudaEventDisableTiming));
    }

    const bool f32 = c.precision == NB_PREC_FP32;
    e->unit_bodies = f32 ? 4u : 1u;
    e->unit_bytes = f32 ? 64 : 32;
    e->vel_unit_bytes = f32 ? 48 : 32;
    // slices: equal unit counts per rank (in-place all-gather), on j-tile boundaries so the
    // force sum can be cut into "own slice" and "gathered" tiles
    const size_t units = (e->n + e->unit_bodies - 1) / e->unit_bodies;
    size_t per = (units + c.world - 1) / c.world;
    // FP32: the large size class cuts j into tiles of 128 quads (nb_direct_f32.cu); slices long
    // enough to get it are whole tiles of that size, so the own-slice / gathered split stays possible
    const size_t align = f32 ? (per >= 2048 ? 2 * kSliceAlignF32 : kSliceAlignF32) : kSliceAlignF64;
    if (c.world > 1) per = (per + align - 1) / align * align;
    if (per == 0) per = 1;
    if (per * (size_t)c.world * e->unit_bodies > 0x7fffffffu) {
        set_err("nb_init", "too many bodies for 32-bit indices");
        return NB_E_INVALID;
    }
    e->slice_units = (unsigned)per;
    e->n_units = (unsigned)(per * c.world);
    e->i_begin = (unsigned)(per * c.rank);
    e->i_end = (unsigned)(per * (c.rank + 1));
    // units past the last real body need no force evaluation
    const unsigned real_units = (unsigned)units;
    if (e->i_end > real_units) e->i_end = real_units > e->i_begin ? real_units : e->i_begin;

    const unsigned my_units = e->i_end - e->i_begin;
    const SliceGeom sg{my_units ? my_units : 1, e->n_units, (unsigned)(per * c.rank), (unsigned)per};
    cudaError_t pe = f32 ? plan_direct_f32(&e->plan, sg, e->sm_count, c.tune_ctas_per_sm, c.tune_flags)
                         : plan_direct_f64(&e->plan, sg, e->sm_count, c.tune_ctas_per_sm, c.tune_flags);
    NB_CUDA(pe);

    const size_t qbytes = (size_t)e->n_units * e->unit_bytes;
    const size_t vbytes = (size_t)e->n_units * e->vel_unit_bytes;
    NB_CUDA(cudaMalloc(&e->q[0], qbytes));
    NB_CUDA(cudaMalloc(&e->q[1], qbytes));
    NB_CUDA(cudaMalloc(&e->vel, vbytes));
    NB_CUDA(cudaMalloc(&e->acc, vbytes));
    NB_CUDA(cudaMemsetAsync(e->acc, 0, vbytes, e->stream));
    NB_CUDA(cudaMalloc(&e->workspace, e->plan.workspace_bytes ? e->plan.workspace_bytes : 16));
    if (e->plan.partial_bytes) NB_CUDA(cudaMalloc(&e->partial, e->plan.partial_bytes));
    e->bh = c.algorithm == NB_ALGO_BARNES_HUT;
    e->count_visits = (c.tune_flags & NB_TUNE_COUNT_VISITS) != 0;
    e->bh_graphs_off = (c.tune_flags & NB_TUNE_NO_GRAPH) != 0;
    if (e->bh && bh_node_cap_of(e->n) > (unsigned long long)BH_CODE_ID) {
        // node ids share a 32-bit stack word with the child count and the bucket bit
        set_err("nb_init", "too many bodies for the Barnes-Hut node ids (27 bits)");
        return NB_E_INVALID;
    }
    e->bh_pipe = e->bh && c.world > 1 && e->n > 0 && !(c.tune_flags & NB_TUNE_REPLICATED_SORT);
    if (e->bh && c.world > 1 && e->n > 1 && !(c.tune_flags & NB_TUNE_NO_REORDER)) {
        const char *ev = getenv("NB_BH_RELAYOUT_EVERY"); // same value on every rank (the call is collective)
        e->relayout_every = ev ? atoi(ev) : 32;
    }
    e->bh_emulate_runs = (e->bh && c.world == 1) ? ((c.tune_flags >> 24) & 0xf) : 0;
    if (e->bh) {
        size_t n_keys = e->n;
        if (c.world > 1) n_keys = (size_t)e->slice_units * c.world;
        if (e->bh_emulate_runs > 1) {
            const size_t per = ((e->n + e->bh_emulate_runs - 1) / e->bh_emulate_runs + 127) / 128 * 128;
            n_keys = per * e->bh_emulate_runs;
        }
        NB_CUDA(bh_alloc(&e->tree, (unsigned)e->n, (unsigned)n_keys, (unsigned)c.world));
    }
    if (e->bh && e->n > 1 && !(c.tune_flags & NB_TUNE_NO_REORDER)) {
        NB_CUDA(cudaMalloc(&e->user_of, e->n * sizeof(unsigned)));
        NB_CUDA(cudaMalloc(&e->order_tmp, e->n * 4 * sizeof(double)));
        if (const char *dbg = getenv("NB_DEBUG_OWN_SLICE")) {
            unsigned w = 0, r = 0;
            if (c.world == 1 && sscanf(dbg, "%u:%u", &w, &r) == 2 && w > 1 && r < w) {
                e->dbg_world = w;
                e->dbg_rank = r;
                e->dbg_len = (unsigned)(((e->n + w - 1) / w + kSliceAlignF64 - 1) / kSliceAlignF64 * kSliceAlignF64);
            }
        }
        if ((c.world > 1 || e->dbg_world) && !(c.tune_flags & NB_TUNE_NO_INTERLEAVE)) {
            NB_CUDA(cudaMalloc(&e->morton_tmp, e->n * sizeof(unsigned)));
            NB_CUDA(cudaMalloc(&e->chunk_base, (e->n / kDealChunk + 1) * sizeof(unsigned)));
        }
    }
    NB_CUDA(cudaMalloc(&e->counters, e->plan.counter_bytes ? e->plan.counter_bytes : 16));
    NB_CUDA(cudaMemsetAsync(e->counters, 0, e->plan.counter_bytes ? e->plan.counter_bytes : 16,
                            e->stream));
    e->io_per = (e->n + (size_t)c.world - 1) / (size_t)c.world;
    const size_t stage_rows = e->io_per * (size_t)c.world + 1; // whole io slices of every rank
    NB_CUDA(cudaMalloc(&e->stage_pos, stage_rows * 4 * sizeof(double)));
    NB_CUDA(cudaMalloc(&e->stage_vel, stage_rows * 3 * sizeof(double)));
    NB_CUDA(cudaMalloc(&e->mass_cls, 4 * sizeof(unsigned)));
    NB_CUDA(cudaMalloc(&e->energy_part, (size_t)(energy_blocks((unsigned)e->n) + 1) * 2 * sizeof(double)));

    if (c.world > 1) {
        rc = load_nccl();
        if (rc) return rc;
        if (!c.comm_id) {
            set_err("nb_init", "world > 1 needs cfg.comm_id from nb_comm_id()");
            return NB_E_INVALID;
        }
        nccl_id_t id;
        memcpy(&id, c.comm_id, NB_COMM_ID_BYTES);
        NB_NCCL(g_nccl.CommInitRank(&e->comm, c.world, id, c.rank));
    }
    e->cfg.comm_id = nullptr; // caller's buffer is not kept

    rc = upload(e, pos, vel, false);
    if (rc) return rc;
    e->st.n = e->n;
    e->st.n_padded = (size_t)e->n_units * e->unit_bodies;
    e->st.slice_begin = (size_t)e->i_begin * e->unit_bodies;
    e->st.slice_end = (size_t)e->i_end * e->unit_bodies;
    if (e->st.slice_end > e->n) e->st.slice_end = e->n;
    if (e->st.slice_begin > e->n) e->st.slice_begin = e->n;
    e->st.grid = (int)e->plan.geom[NB_GEOM_FULL].grid;
    e->st.block = (int)e->plan.block;
    return NB_OK;
}

int nb_init(nb_engine **out, const nb_config *cfg, size_t n, const double *pos_xyzm,
            const double *vel_xyz)
{
    if (!out || !cfg) return NB_E_INVALID;
    *out = nullptr;
    if ((n && !pos_xyzm) || cfg->world < 1 || cfg->rank < 0 || cfg->rank >= cfg->world ||
        (cfg->precision != NB_PREC_FP64 && cfg->precision != NB_PREC_FP32) ||
        !(cfg->dt == cfg->dt) || cfg->eps2 < 0) {
        set_err("nb_init", "bad configuration");
        return NB_E_INVALID;
    }
    if (cfg->algorithm != NB_ALGO_DIRECT && cfg->algorithm != NB_ALGO_BARNES_HUT) {
        set_err("nb_init", "unknown algorithm");
        return NB_E_INVALID;
    }
    if (cfg->algorithm == NB_ALGO_DIRECT && !(cfg->eps2 > 0)) {
        // the direct kernels include the j = i term like the reference (serial/exhaustive.c:163)
        // and rely on the softening to make it exactly 0; with eps2 = 0 it is 0 * inf = NaN
        set_err("nb_init", "the direct sum needs eps2 > 0 (the j = i term is zeroed by the softening)");
        return NB_E_INVALID;
    }
    if (cfg->algorithm == NB_ALGO_BARNES_HUT && cfg->precision != NB_PREC_FP64) {
        set_err("nb_init", "Barnes-Hut runs in FP64 (the reference's precision) only");
        return NB_E_INVALID;
    }
    if (cfg->algorithm == NB_ALGO_BARNES_HUT && !(cfg->theta >= 0)) {
        set_err("nb_init", "bad theta");
        return NB_E_INVALID;
    }
    int ndev = nb_device_count();
    if (ndev <= 0 || cfg->device < 0 || cfg->device >= ndev) {
        set_err("nb_init", "no usable CUDA device; the engine has no CPU fallback");
        return NB_E_NODEVICE;
    }
    nb_engine *e = new (std::nothrow) nb_engine();
    if (!e) return NB_E_NOMEM;
    memset(&e->st, 0, sizeof(e->st));
    e->cfg = *cfg;
    e->n = n;
    int rc = init_impl(e, pos_xyzm, vel_xyz);
    if (rc) {
        nb_free(e);
        return rc;
    }
    *out = e;
    return NB_OK;
}

int nb_upload(nb_engine *e, const double *pos_xyzm, const double *vel_xyz)
{
    if (!e || (e->n && !pos_xyzm)) return NB_E_INVALID;
    int rc = set_device(e);
    if (rc) return rc;
    return upload(e, pos_xyzm, vel_xyz, false);
}

int nb_io_slice(nb_engine *e, size_t *begin, size_t *count)
{
    if (!e || !begin || !count) return NB_E_INVALID;
    io_rows(e, begin, count);
    return NB_OK;
}

int nb_upload_slice(nb_engine *e, const double *pos_rows, const double *vel_rows)
{
    if (!e) return NB_E_INVALID;
    size_t b, c;
    io_rows(e, &b, &c);
    if (c && !pos_rows) return NB_E_INVALID;
    int rc = set_device(e);
    if (rc) return rc;
    return upload(e, pos_rows, vel_rows, true);
}

int nb_step(nb_engine *e, int n_steps)
{
    if (!e || n_steps < 0) return NB_E_INVALID;
    int rc = set_device(e);
    if (rc) return rc;
    NB_CUDA(cudaEventRecord(e->ev_step0, e->stream));
    struct Restore { nb_engine *e; ~Restore() { e->time_force = true; } } restore{e};
    for (int s = 0; s < n_steps; ++s) {
        e->time_force = (s == n_steps - 1);
        if (e->relayout_every > 0 && e->state == MIDSTEP && e->steps_since_layout >= e->relayout_every) {
            rc = relayout(e);
            if (rc) return rc;
        }
        ++e->steps_since_layout;
        if (e->state == SYNCED_ACC) {
            NB_CUDA(kickdrift(e->cfg.precision, e->q[e->cur], e->q[e->cur ^ 1], e->vel, e->acc,
                              e->i_begin, e->i_end, (unsigned)e->n, e->cfg.dt, e->stream));
            e->st.kernel_launches++;
            e->vel_complete = (e->cfg.world <= 1);
            rc = exchange_positions(e);
            if (rc) return rc;
        } else if (e->state == SYNCED_NOACC) {
            rc = launch_force(e, NB_F_KICK1_DRIFT);
            if (rc) return rc;
        } else {
            rc = launch_force(e, NB_F_KICK2 | NB_F_KICK1_DRIFT);
            if (rc) return rc;
        }
        e->state = MIDSTEP;
        e->st.steps_done++;
    }
    NB_CUDA(cudaEventRecord(e->ev_step1, e->stream));
    e->have_step_ev = true;
    e->step_ended_in_gather = e->gather_pending && n_steps > 0;
    return NB_OK;
}

int nb_sync(nb_engine *e)
{
    if (!e) return NB_E_INVALID;
    int rc = set_device(e);
    if (rc) return rc;
    NB_CUDA(cudaStreamSynchronize(e->stream));
    if (e->comm_stream) NB_CUDA(cudaStreamSynchronize(e->comm_stream));
    if (e->copy_stream) NB_CUDA(cudaStreamSynchronize(e->copy_stream));
    return NB_OK;
}

int nb_frame_begin(nb_engine *e, double *pos_xyzm)
{
    if (!e || !pos_xyzm) return NB_E_INVALID;
    int rc = set_device(e);
    if (rc) return rc;
    const unsigned n = (unsigned)e->n;
    if (!e->copy_stream) {
        NB_CUDA(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
        for (int k = 0; k < 2; ++k) {
            NB_CUDA(cudaMalloc(&e->frame_stage[k], (size_t)(n ? n : 1) * 4 * sizeof(double)));
            NB_CUDA(cudaEventCreateWithFlags(&e->ev_frame_ready[k], cudaEventDisableTiming));
            NB_CUDA(cudaEventCreateWithFlags(&e->ev_frame_done[k], cudaEventDisableTiming));
        }
    }
    rc = wait_gather(e); // every rank's positions of this step are in place
    if (rc) return rc;
    const int k = (int)(e->frame_seq++ & 1u);
    // the slot's previous snapshot must have left the device before it is overwritten
    if (e->frame_host[k]) NB_CUDA(cudaStreamWaitEvent(e->stream, e->ev_frame_done[k], 0));
    NB_CUDA(unpack_bodies(e->cfg.precision, e->q[e->cur], n, 4, e->frame_stage[k], e->user_of, e->stream));
    e->st.kernel_launches++;
    NB_CUDA(cudaEventRecord(e->ev_frame_ready[k], e->stream));
    NB_CUDA(cudaStreamWaitEvent(e->copy_stream, e->ev_frame_ready[k], 0));
    if (n)
        NB_CUDA(cudaMemcpyAsync(pos_xyzm, e->frame_stage[k], (size_t)n * 4 * sizeof(double),
                                cudaMemcpyDeviceToHost, e->copy_stream));
    NB_CUDA(cudaEventRecord(e->ev_frame_done[k], e->copy_stream));
    e->frame_host[k] = pos_xyzm;
    return NB_OK;
}

int nb_frame_wait(nb_engine *e, const double *pos_xyzm)
{
    if (!e) return NB_E_INVALID;
    int rc = set_device(e);
    if (rc) return rc;
    if (!e->copy_stream) return NB_OK; // nothing was ever started
    bool known = false;
    for (int k = 0; k < 2; ++k)
        if (pos_xyzm && e->frame_host[k] == pos_xyzm) {
            NB_CUDA(cudaEventSynchronize(e->ev_frame_done[k]));
            known = true;
        }
    if (!known) NB_CUDA(cudaStreamSynchronize(e->copy_stream)); // NULL or an older snapshot
    return NB_OK;
}

int nb_host_alloc(void **p, size_t bytes)
{
    if (!p) return NB_E_INVALID;
    *p = nullptr;
    NB_CUDA(cudaHostAlloc(p, bytes ? bytes : 1, cudaHostAllocDefault));
    return NB_OK;
}

int nb_host_free(void *p)
{
    if (p) NB_CUDA(cudaFreeHost(p));
    return NB_OK;
}

// rows [row0, row0 + rows) of the caller's order go to the host arrays (which hold just those rows)
static int gather_rows(nb_engine *e, double *pos_xyzm, double *vel_xyz, size_t row0, size_t rows)
{
    int rc = set_device(e);
    if (rc) return rc;
    const unsigned n = (unsigned)e->n;
    if (vel_xyz) {
        rc = close_step(e);
        if (rc) return rc;
        rc = ensure_vel_complete(e);
        if (rc) return rc;
        NB_CUDA(unpack_bodies(e->cfg.precision, e->vel, n, 3, e->stage_vel, e->user_of, e->stream));
        e->st.kernel_launches++;
        if (rows)
            NB_CUDA(cudaMemcpyAsync(vel_xyz, e->stage_vel + 3 * row0, rows * 3 * sizeof(double),
                                    cudaMemcpyDeviceToHost, e->stream));
    }
    if (pos_xyzm) {
        rc = wait_gather(e);
        if (rc) return rc;
        NB_CUDA(unpack_bodies(e->cfg.precision, e->q[e->cur], n, 4, e->stage_pos, e->user_of, e->stream));
        e->st.kernel_launches++;
        if (rows)
            NB_CUDA(cudaMemcpyAsync(pos_xyzm, e->stage_pos + 4 * row0, rows * 4 * sizeof(double),
                                    cudaMemcpyDeviceToHost, e->stream));
    }
    NB_CUDA(cudaStreamSynchronize(e->stream));
    if (e->bh && e->st.force_launches) { // a truncated tree must not pass for a result
        nb_tree_info ti;
        rc = nb_tree_stats(e, &ti);
        if (rc) return rc;
    }
    return NB_OK;
}

int nb_gather(nb_engine *e, double *pos_xyzm, double *vel_xyz)
{
    if (!e) return NB_E_INVALID;
    return gather_rows(e, pos_xyzm, vel_xyz, 0, e->n);
}

int nb_gather_slice(nb_engine *e, double *pos_rows, double *vel_rows)
{
    if (!e) return NB_E_INVALID;
    size_t b, c;
    io_rows(e, &b, &c);
    return gather_rows(e, pos_rows, vel_rows, b, c);
}

int nb_accelerations(nb_engine *e, double *acc_xyz)
{
    if (!e || !acc_xyz) return NB_E_INVALID;
    int rc = set_device(e);
    if (rc) return rc;
    if (e->state == MIDSTEP) {
        rc = close_step(e);
    } else if (e->state == SYNCED_NOACC) {
        rc = launch_force(e, NB_F_STORE_ACC);
        if (!rc) e->state = SYNCED_ACC;
    }
    if (rc) return rc;
    rc = all_gather_units(e, e->acc, e->vel_unit_bytes);
    if (rc) return rc;
    rc = wait_gather(e);
    if (rc) return rc;
    const unsigned n = (unsigned)e->n;
    NB_CUDA(unpack_bodies(e->cfg.precision, e->acc, n, 3, e->stage_vel, e->user_of, e->stream));
    e->st.kernel_launches++;
    if (n)
        NB_CUDA(cudaMemcpyAsync(acc_xyz, e->stage_vel, (size_t)n * 3 * sizeof(double),
                                cudaMemcpyDeviceToHost, e->stream));
    NB_CUDA(cudaStreamSynchronize(e->stream));
    return NB_OK;
}

int nb_energy(nb_engine *e, double *KE, double *PE)
{
    if (!e || !KE || !PE) return NB_E_INVALID;
    int rc = set_device(e);
    if (rc) return rc;
    rc = close_step(e);
    if (rc) return rc;
    rc = ensure_vel_complete(e);
    if (rc) return rc;
    rc = wait_gather(e);
    if (rc) return rc;
    const unsigned n = (unsigned)e->n;
    NB_CUDA(unpack_bodies(e->cfg.precision, e->vel, n, 3, e->stage_vel, nullptr, e->stream));
    NB_CUDA(unpack_bodies(e->cfg.precision, e->q[e->cur], n, 4, e->stage_pos, nullptr, e->stream)); // any order
    NB_CUDA(energy_partials(e->stage_pos, e->stage_vel, n, e->energy_part, e->stream));
    e->st.kernel_launches += 3;
    const unsigned nb = energy_blocks(n);
    double *h = (double *)malloc((size_t)(nb + 1) * 2 * sizeof(double));
    if (!h) return NB_E_NOMEM;
    if (nb)
        NB_CUDA(cudaMemcpyAsync(h, e->energy_part, (size_t)nb * 2 * sizeof(double),
                                cudaMemcpyDeviceToHost, e->stream));
    NB_CUDA(cudaStreamSynchronize(e->stream));
    double ke = 0, pe = 0;
    for (unsigned b = 0; b < nb; ++b) {
        ke += h[2 * b];
        pe += h[2 * b + 1];
    }
    free(h);
    *KE = 0.5 * ke;
    *PE = -e->cfg.G * pe;
    return NB_OK;
}

int nb_energy_static(int device, const double *d_pos_xyzm, const double *d_vel_xyz, size_t n,
                     double G, double *d_ke_pe, void *stream)
{
    if (!d_ke_pe || (n && (!d_pos_xyzm || !d_vel_xyz))) {
        set_err("nb_energy_static", "null pointer");
        return NB_E_INVALID;
    }
    if (n > (size_t)NB_ENERGY_STATIC_MAX_BODIES) {
        set_err("nb_energy_static", "n exceeds NB_ENERGY_STATIC_MAX_BODIES (1048576): the static scratch array holds no more block partials");
        return NB_E_INVALID;
    }
    NB_CUDA(energy_static(device, d_pos_xyzm, d_vel_xyz, (unsigned)n, G, d_ke_pe,
                          static_cast<cudaStream_t>(stream)));
    return NB_OK;
}

int nb_tree_stats(nb_engine *e, nb_tree_info *out)
{
    if (!e || !out) return NB_E_INVALID;
    if (!e->bh) {
        set_err("nb_tree_stats", "engine was not created with NB_ALGO_BARNES_HUT");
        return NB_E_STATE;
    }
    int rc = set_device(e);
    if (rc) return rc;
    BhStats bs;
    NB_CUDA(bh_stats(&e->tree, &bs, e->stream));
    out->nodes = bs.nodes;
    out->depth = bs.depth;
    out->root_border = bs.root_border;
    out->interactions = bs.interactions;
    out->node_visits = bs.visits;
    out->child_tests = bs.child_tests;
    for (int k = 0; k < 8; ++k) out->stage_ms[k] = k < BH_STAGES ? bs.stage_ms[k] : 0.0;
    if (e->have_x_ev && e->comm_stream) { // the last position exchange, waiting for the slowest rank included
        NB_CUDA(cudaStreamSynchronize(e->comm_stream));
        float ms = 0.f;
        NB_CUDA(cudaEventElapsedTime(&ms, e->ev_x0, e->ev_gather));
        out->stage_ms[7] = ms;
    }
    out->pool_overflow = (int)bs.pool_overflow;
    out->stack_overflow = (int)bs.stack_overflow;
    if (bs.pool_overflow || bs.stack_overflow) {
        set_err("nb_tree_stats", bs.pool_overflow ? "octree node pool exhausted" : "traversal stack overflow");
        return NB_E_NOMEM;
    }
    return NB_OK;
}

int nb_get_stats(nb_engine *e, nb_stats *out)
{
    if (!e || !out) return NB_E_INVALID;
    int rc = set_device(e);
    if (rc) return rc;
    NB_CUDA(cudaStreamSynchronize(e->stream));
    if (e->comm_stream) NB_CUDA(cudaStreamSynchronize(e->comm_stream));
    float ms = 0.f;
    if (e->have_step_ev) {
        NB_CUDA(cudaEventElapsedTime(&ms, e->ev_step0, e->ev_step1));
        e->st.last_step_ms = ms;
        if (e->step_ended_in_gather) { // the step is over when the exchange is
            NB_CUDA(cudaEventElapsedTime(&ms, e->ev_step0, e->ev_gather));
            if (ms > e->st.last_step_ms) e->st.last_step_ms = ms;
        }
    }
    if (e->have_force_ev) {
        NB_CUDA(cudaEventElapsedTime(&ms, e->ev_f0, e->ev_f1));
        e->st.last_force_ms = ms;
    }
    e->st.uniform_mass = e->uniform ? 1 : 0;
    e->st.kernel_variant = e->bh ? -1 : e->plan.variant;
    *out = e->st;
    return NB_OK;
}

} // extern "C"
