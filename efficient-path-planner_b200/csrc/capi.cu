// capi.cu — the C ABI of libepp_b200.so (include/epp_c_api.h): world epochs on the device,
// host-buffer staging, and dispatch into the kernels.  No CPU fallback exists anywhere in
// this file: a missing device or CUDA failure is reported, never papered over.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <limits>
#include <memory>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/epp_c_api.h"
#include "kernels.h"
#include "world_host.h"

using namespace epp;

namespace {

thread_local std::string tl_error;

int fail(int status, const std::string &detail = std::string())
{
    tl_error = status_message(status);
    if (!detail.empty()) tl_error += ": " + detail;
    return status;
}

int cuda_fail(cudaError_t e, const char *what)
{
    if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver)
        return fail(EPP_ERR_NO_DEVICE, std::string(what) + ": " + cudaGetErrorString(e));
    return fail(EPP_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

#define EPP_CUDA(call)                                         \
    do {                                                       \
        cudaError_t e__ = (call);                              \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call);  \
    } while (0)

// Tuning knobs (epp_set_tuning / epp_get_tuning): one table, looked up by name.
struct Knob {
    const char *name;
    std::atomic<int> value;
};
Knob g_knobs[] = {
    {"states_blocks_per_sm", {8}},       // direct form of K1: persistent blocks per SM
    {"segments_blocks_per_sm", {4}},     // direct forms of K2: persistent blocks per SM
    {"segments_variant", {0}},           // 0: auto, 1: direct form with records in global memory, 2: warp-pipelined, 3: direct forms only
    {"segments_warp_min_n", {32768}},    // K2: batches from this size up take the warp-pipelined form
    {"segments_warp_threads", {1024}},   // K2 warp form: threads per block (1 producer warp + consumers)
    {"segments_warp_slots", {0}},        // K2 warp form: ring slots (0: all that fit, up to 64)
    {"segments_block", {0}},             // k_segments_rs threads per block: 256, 512, 1024; 0 = by batch size
    {"traj_variant", {0}},               // pre-sampled trajectories: 0 auto (warp-pipelined), 1 direct form only
    {"states_variant", {0}},             // 0: auto, 1: direct, 3: warp-pipelined
    {"states_per_thread", {0}},          // warp form: mask words per chunk (4 default, 8)
    {"states_warp_blocks_per_sm", {1}},  // K1 warp form: resident blocks per SM
    {"states_warp_threads", {1024}},     // K1 warp form: threads per block (1 producer warp + consumers)
    {"states_warp_log2_slots", {0}},     // K1 warp form: ring slots (0: all that fit)
    {"states_warp_min_n", {4096}},       // K1: batches from this size up take the warp-pipelined form
    {"states_rec_smem", {1}},            // K1 warp form: folded OBB records in shared memory for small worlds
    {"states_rec_smem_max", {0}},        // 0: 256 OBBs
    {"states_bits1", {1}},               // worlds of <= 32 OBBs: one-word bin rows (PointWorld<kBpBits1>)
    {"states_pdl", {1}},                 // programmatic dependent launch for the warp forms of K1 / K2
    {"states_late_wait", {1}},           // multi-batch calls: batches after the first do not wait for the previous launch before streaming
    {"states_multi_blocks_per_sm", {2}}, // multi-batch state calls: resident blocks per SM (threads per block = 1024 / this)
    {"segments_multi_blocks_per_sm", {2}}, // multi-batch segment calls: resident blocks per SM (threads per block = 1024 / this; 1: one 1024-thread block)
    {"states_multi_grid_per_sm", {1}},   // multi-batch calls: blocks per SM in each launch's grid (0: all that are resident)
    {"states_debug", {0}},               // profiling builds only
    {"poly_traj_per_warp", {0}},         // K3b: trajectories per block, 0 = by batch size
    {"poly_balance", {1}},               // K3b: trajectories per block of big batches chosen to balance the SMs (0: round-2 halving rule)
    {"poly_ring", {0}},                  // K3b: rounds of parked samples per block (2 .. 8; 0 = by variant)
    {"poly_blocks_per_sm", {0}},         // K3b: 4 = 256-thread blocks of 64-register threads, 5 = never those, 0 = by batch size
    {"poly_wait", {0}},                  // K3b: 1 = consumers sleep on the round's barrier, 0 = nanosleep poll
    {"poly_threads", {0}},               // K3b: threads per block: 0 = by batch size; 256 (48 registers, 5 per SM), 224, 192, 512
    {"gather_group_chunks", {9}},        // fused exchange: chunks per NVLink run (1: every chunk alone)
    {"gather_bulk_min_words", {65536}},  // exchange: slots from this many words up are copied straight into the peers' buffers (0: never)
    {"gather_debug", {0}},               // profiling only: 1 no wait/unpack, 2 no peer stores
    {"sm_limit", {0}},                   // 0: all SMs; > 0: size grids for this many SMs (several ranks on one GPU in tests)
    {"bits_bins", {0}},                  // 0: by world size
    {"grid_max_cells", {1 << 22}},       // grid broad phase: upper bound on cells; read at commit
    {"grid_occ_kb", {72}},               // shared-memory budget of the grid's occupancy bitmap; read at commit
    {"grid_cell_recs", {1}},             // grid: candidates read from records stored in cell order (0: item -> AabbRec -> ObbRec); read at commit
    {"bits_range_d", {5}},               // bits broad phase, rays: (first bin, span) rows for boxes spanning < d bins on x and y (0: cumulative rows only); read at commit
    {"segments_recs_at_commit", {1}},    // K2: folded records for both canPassGate values built at commit (0: folded per block); read at commit
    {"bits_cmask", {1}},                 // bits broad phase, W > 1: coarse-cell word masks (0: round-1 summaries); read at commit
    {"morton_sort", {1}},                // device records along a Morton curve (worlds of > 32 OBBs); read at commit
    {"single_mailbox", {1}},             // epp_is_valid / epp_check_motion through the resident mailbox warp (0: launch + sync per call)
    {"single_mailbox_idle_us", {1000}},  // the mailbox warp leaves after this long without a request
    {"bits_max_obbs", {1024}},           // worlds above this many OBBs take the grid (auto) / are refused for EPP_BP_BITS; at most 1024
};

std::atomic<int> *knob(const char *name)
{
    for (Knob &k : g_knobs)
        if (std::strcmp(k.name, name) == 0) return &k.value;
    return nullptr;
}

int knob_value(const char *name) { return knob(name)->load(); }

constexpr int kBitsMaxObbs = 1024;

// One immutable committed world on one device.
struct DeviceEpoch {
    int device = -1;
    int sm_count = 148;
    uint64_t epoch = 0;
    int bp = EPP_BP_BITS;
    void *blob = nullptr;
    WorldView view{};
    ~DeviceEpoch()
    {
        if (blob) {
            int prev = -1;
            cudaGetDevice(&prev);
            cudaSetDevice(device);
            cudaFree(blob); // implicit device synchronisation: in-flight kernels finish first
            if (prev >= 0) cudaSetDevice(prev);
        }
    }
};

} // namespace

struct epp_world {
    HostWorld host;
    mutable std::mutex mu;
    std::shared_ptr<DeviceEpoch> current;
    bool dirty = true;
    uint64_t epoch = 0;
    int bp_request = EPP_BP_AUTO;
    epp_world(double g, double o) : host(g, o) {}
};

namespace {

// Per-thread, per-device staging: a stream, pinned host memory and device scratch.  Gives
// the re-entrancy the reference's concurrent planner threads need without a global lock.
// Host side of the single-unit mailbox (k_mailbox): one per host thread and device, serving one world epoch at a time.
struct Mailbox {
    MailboxSlot *slot = nullptr;          // mapped pinned memory (host pointer == device pointer under UVA)
    cudaStream_t stream = nullptr;        // the resident warp's own stream
    std::shared_ptr<DeviceEpoch> epoch;   // keeps the blob the warp reads alive
    uint32_t seq = 0;
    bool launched = false;                // a warp was launched and has not been seen to exit
    void quit();                          // ask the resident warp to leave and wait until it has
    ~Mailbox();
};

struct Staging {
    int device = -1;
    cudaStream_t stream = nullptr;
    unsigned char *h = nullptr;
    size_t h_cap = 0;
    unsigned char *d = nullptr;
    size_t d_cap = 0;
    Mailbox mb;
    ~Staging()
    {
        if (device < 0) return;
        mb.quit();
        // Thread exit may happen after the primary context is gone; ignore errors.
        if (stream) cudaStreamDestroy(stream);
        if (h) cudaFreeHost(h);
        if (d) cudaFree(d);
    }
};

thread_local std::vector<std::unique_ptr<Staging>> tl_staging;

int get_staging(int device, Staging **out)
{
    for (auto &s : tl_staging)
        if (s->device == device) {
            *out = s.get();
            return EPP_OK;
        }
    std::unique_ptr<Staging> s(new Staging);
    EPP_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    s->device = device;
    *out = s.get();
    tl_staging.push_back(std::move(s));
    return EPP_OK;
}

int reserve_host(Staging *s, size_t bytes)
{
    if (bytes <= s->h_cap) return EPP_OK;
    EPP_CUDA(cudaStreamSynchronize(s->stream));
    if (s->h) cudaFreeHost(s->h);
    s->h = nullptr;
    s->h_cap = 0;
    size_t cap = std::max<size_t>(bytes + bytes / 2, 1 << 16);
    EPP_CUDA(cudaHostAlloc((void **)&s->h, cap, cudaHostAllocMapped));
    s->h_cap = cap;
    return EPP_OK;
}

int reserve_device(Staging *s, size_t bytes)
{
    if (bytes <= s->d_cap) return EPP_OK;
    EPP_CUDA(cudaStreamSynchronize(s->stream));
    if (s->d) cudaFree(s->d);
    s->d = nullptr;
    s->d_cap = 0;
    size_t cap = std::max<size_t>(bytes + bytes / 2, 1 << 20);
    EPP_CUDA(cudaMalloc((void **)&s->d, cap));
    s->d_cap = cap;
    return EPP_OK;
}

bool is_pinned(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

constexpr size_t kChunk = 8u << 20; // staging chunk for pageable host memory

// H2D from arbitrary host memory: pinned sources go straight to the copy engine, pageable
// ones are bounced through two pinned chunks so memcpy overlaps the DMA.
int upload(Staging *s, void *dst, const void *src, size_t bytes, size_t bounce_off)
{
    if (bytes == 0) return EPP_OK;
    if (is_pinned(src)) {
        EPP_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s->stream));
        return EPP_OK;
    }
    EPP_CUDA(cudaStreamSynchronize(s->stream)); // bounce chunks may still feed an earlier copy
    cudaEvent_t ev[2];
    EPP_CUDA(cudaEventCreateWithFlags(&ev[0], cudaEventDisableTiming));
    EPP_CUDA(cudaEventCreateWithFlags(&ev[1], cudaEventDisableTiming));
    int rc = EPP_OK;
    size_t off = 0;
    for (int i = 0; off < bytes; ++i) {
        const size_t len = std::min(kChunk, bytes - off);
        unsigned char *b = s->h + bounce_off + (size_t)(i & 1) * kChunk;
        if (i >= 2) cudaEventSynchronize(ev[i & 1]);
        std::memcpy(b, (const unsigned char *)src + off, len);
        cudaError_t e = cudaMemcpyAsync((unsigned char *)dst + off, b, len, cudaMemcpyHostToDevice, s->stream);
        if (e == cudaSuccess) e = cudaEventRecord(ev[i & 1], s->stream);
        if (e != cudaSuccess) {
            rc = cuda_fail(e, "upload");
            break;
        }
        off += len;
    }
    cudaEventDestroy(ev[0]);
    cudaEventDestroy(ev[1]);
    return rc;
}

int snapshot(epp_world *w, std::shared_ptr<DeviceEpoch> &ep)
{
    if (!w) return fail(EPP_ERR_INVALID_ARG, "null world");
    std::lock_guard<std::mutex> lk(w->mu);
    // Mutations that have not been committed yet are invisible: a check sees the world as of the last commit (another
    // thread may be between its update and its commit, src/OnlineTrajGenerator.cpp:146 vs :324-340).
    if (!w->current) return fail(EPP_ERR_NOT_COMMITTED);
    ep = w->current;
    return EPP_OK;
}

LaunchCfg launch_cfg(const DeviceEpoch &ep)
{
    LaunchCfg c;
    c.sm_count = ep.sm_count;
    const int sm_limit = knob_value("sm_limit");
    if (sm_limit > 0 && sm_limit < c.sm_count) c.sm_count = sm_limit;
    c.states_blocks_per_sm = knob_value("states_blocks_per_sm");
    c.segments_blocks_per_sm = knob_value("segments_blocks_per_sm");
    c.segments_variant = knob_value("segments_variant");
    c.segments_warp_min_n = knob_value("segments_warp_min_n");
    c.segments_warp_threads = knob_value("segments_warp_threads");
    c.segments_warp_slots = knob_value("segments_warp_slots");
    c.traj_variant = knob_value("traj_variant");
    c.segments_block = knob_value("segments_block");
    c.states_variant = knob_value("states_variant");
    c.states_per_thread = knob_value("states_per_thread");
    c.states_warp_blocks_per_sm = knob_value("states_warp_blocks_per_sm");
    c.states_warp_threads = knob_value("states_warp_threads");
    c.states_warp_log2_slots = knob_value("states_warp_log2_slots");
    c.states_warp_min_n = knob_value("states_warp_min_n");
    c.states_rec_smem = knob_value("states_rec_smem");
    c.states_rec_smem_max = knob_value("states_rec_smem_max");
    c.states_bits1 = knob_value("states_bits1");
    c.states_pdl = knob_value("states_pdl");
    c.states_debug = knob_value("states_debug");
    c.poly_traj_per_warp = knob_value("poly_traj_per_warp");
    c.poly_threads = knob_value("poly_threads");
    c.poly_balance = knob_value("poly_balance");
    c.poly_ring = knob_value("poly_ring");
    c.poly_wait = knob_value("poly_wait");
    c.poly_blocks_per_sm = knob_value("poly_blocks_per_sm");
    c.gather_debug = knob_value("gather_debug");
    c.gather_group_chunks = knob_value("gather_group_chunks");
    c.gather_bulk_min_words = knob_value("gather_bulk_min_words");
    return c;
}

// RAII: run on the epoch's device, restore the caller's device afterwards.
struct DeviceGuard {
    int prev = -1;
    bool switched = false;
    cudaError_t enter(int device)
    {
        cudaError_t e = cudaGetDevice(&prev);
        if (e != cudaSuccess) return e;
        if (prev != device) {
            e = cudaSetDevice(device);
            switched = (e == cudaSuccess);
        }
        return e;
    }
    ~DeviceGuard()
    {
        if (switched) cudaSetDevice(prev);
    }
};

size_t mask_words(size_t n) { return (n + 31) / 32; }

// Common body of the host-buffer point/segment checks.
int check_host(epp_world *w, const double *a, const double *b, size_t n, int kind /*0 point,1 ray*/, int mode,
               int can_pass_gate, double min_distance, uint32_t *mask)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n == 0) return EPP_OK;
    if (!a || !mask || (kind == 1 && !b)) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    DeviceGuard guard;
    EPP_CUDA(guard.enter(ep->device));
    Staging *s;
    rc = get_staging(ep->device, &s);
    if (rc) return rc;

    const size_t in_bytes = n * 3 * sizeof(double);
    const size_t words = mask_words(n);
    const size_t off_b = align_up(in_bytes, 256);
    const size_t off_mask = kind == 1 ? off_b + align_up(in_bytes, 256) : off_b;
    rc = reserve_device(s, off_mask + words * 4);
    if (rc) return rc;
    const bool mask_pinned = is_pinned(mask);
    // pinned layout: [bounce 2 x kChunk][mask]
    rc = reserve_host(s, 2 * kChunk + (mask_pinned ? 0 : words * 4));
    if (rc) return rc;

    rc = upload(s, s->d, a, in_bytes, 0);
    if (rc) return rc;
    if (kind == 1) {
        rc = upload(s, s->d + off_b, b, in_bytes, 0);
        if (rc) return rc;
    }
    uint32_t *d_mask = (uint32_t *)(s->d + off_mask);
    const LaunchCfg cfg = launch_cfg(*ep);
    cudaError_t e;
    if (kind == 0)
        e = launch_states(cfg, ep->view, ep->bp, (const double *)s->d, n, mode, can_pass_gate, min_distance, d_mask,
                          s->stream);
    else
        e = launch_segments(cfg, ep->view, ep->bp, (const double *)s->d, (const double *)(s->d + off_b), n,
                            can_pass_gate, d_mask, s->stream);
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    uint32_t *h_mask = mask_pinned ? mask : (uint32_t *)(s->h + 2 * kChunk);
    EPP_CUDA(cudaMemcpyAsync(h_mask, d_mask, words * 4, cudaMemcpyDeviceToHost, s->stream));
    EPP_CUDA(cudaStreamSynchronize(s->stream));
    if (!mask_pinned) std::memcpy(mask, h_mask, words * 4);
    return EPP_OK;
}

// ---- single units ------------------------------------------------------------------------------------------------

using Clock = std::chrono::steady_clock;

bool mailbox_exited(const MailboxSlot *slot) { return slot->state == kMailboxExited; }

void Mailbox::quit()
{
    if (!slot) return;
    // At thread / process exit the CUDA context may already be gone: then neither the slot nor the stream may be touched.
    const cudaError_t q = stream ? cudaStreamQuery(stream) : cudaSuccess;
    if (q != cudaSuccess && q != cudaErrorNotReady) {
        cudaGetLastError();
        slot = nullptr;
        stream = nullptr;
        launched = false;
        epoch.reset();
        return;
    }
    if (launched && q == cudaErrorNotReady && !mailbox_exited(slot)) {
        ++seq;
        slot->req[0] = (uint64_t)seq | ((uint64_t)kMailboxQuit << 32);
        std::atomic_thread_fence(std::memory_order_release);
        slot->req[7] = (uint64_t)seq;
        std::atomic_thread_fence(std::memory_order_seq_cst);
        const auto t0 = Clock::now();
        while (!mailbox_exited(slot) && Clock::now() - t0 < std::chrono::milliseconds(200)) {
        }
        if (stream) cudaStreamSynchronize(stream);
    }
    launched = false;
    epoch.reset();
}

Mailbox::~Mailbox()
{
    quit();
    if (stream) cudaStreamDestroy(stream);
    if (slot) cudaFreeHost(slot);
}

// One unit through the resident mailbox warp: no launch and no stream synchronisation per call.
int check_single_mailbox(Staging *s, const std::shared_ptr<DeviceEpoch> &ep, const double *a, const double *b,
                         int can_pass_gate, int *valid)
{
    Mailbox &mb = s->mb;
    if (!mb.slot) {
        EPP_CUDA(cudaHostAlloc((void **)&mb.slot, sizeof(MailboxSlot), cudaHostAllocMapped));
        std::memset((void *)mb.slot, 0, sizeof(MailboxSlot));
        EPP_CUDA(cudaStreamCreateWithFlags(&mb.stream, cudaStreamNonBlocking));
    }
    if (mb.epoch != ep) { // the world was re-committed (or this is the first call): retire the warp of the old epoch
        mb.quit();
        mb.epoch = ep;
    }
    MailboxSlot *slot = mb.slot;
    const uint32_t seq = ++mb.seq;
    const uint32_t flags = (b ? kMailboxSegment : 0u) | (can_pass_gate ? kMailboxCanPassGate : 0u);
    uint64_t w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    std::memcpy(&w[1], a, 24);
    if (b) std::memcpy(&w[4], b, 24);
    for (int i = 1; i < 7; ++i) slot->req[i] = w[i];
    std::atomic_thread_fence(std::memory_order_release); // payload before the sequence numbers (x86: stores stay in order)
    slot->req[0] = (uint64_t)seq | ((uint64_t)flags << 32);
    slot->req[7] = (uint64_t)seq;
    std::atomic_thread_fence(std::memory_order_seq_cst);
    const uint64_t idle_ns = (uint64_t)std::max(1, knob_value("single_mailbox_idle_us")) * 1000ull;
    auto launch = [&]() -> int {
        slot->state = kMailboxRunning;
        std::atomic_thread_fence(std::memory_order_seq_cst);
        cudaError_t e = launch_mailbox(ep->view, ep->bp, slot, seq - 1u, idle_ns, mb.stream);
        if (e != cudaSuccess) return cuda_fail(e, "mailbox launch");
        mb.launched = true;
        return EPP_OK;
    };
    if (!mb.launched || mailbox_exited(slot)) {
        int rc = launch();
        if (rc) return rc;
    }
    const auto t0 = Clock::now();
    uint32_t spins = 0;
    for (;;) {
        if (slot->ack[0] == seq) break;
        if ((++spins & 1023u) == 0u) {
            if (mailbox_exited(slot) && slot->ack[0] != seq) { // the warp left just before this request: start another
                int rc = launch();
                if (rc) return rc;
            }
            if (Clock::now() - t0 > std::chrono::seconds(5)) {
                cudaError_t e = cudaStreamQuery(mb.stream);
                if (e != cudaSuccess && e != cudaErrorNotReady) return cuda_fail(e, "mailbox kernel");
                return fail(EPP_ERR_CUDA, "mailbox: no answer from the device within 5 s");
            }
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    *valid = (int)(slot->ack[1] & 1u);
    return EPP_OK;
}

// One unit through mapped pinned memory: the mailbox warp by default; a single launch + synchronise with the knob off.
int check_single(epp_world *w, const double *a, const double *b, int can_pass_gate, int *valid)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (!a || !valid) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    DeviceGuard guard;
    EPP_CUDA(guard.enter(ep->device));
    Staging *s;
    rc = get_staging(ep->device, &s);
    if (rc) return rc;
    if (knob_value("single_mailbox")) return check_single_mailbox(s, ep, a, b, can_pass_gate, valid);
    rc = reserve_host(s, 256);
    if (rc) return rc;
    double *slot = (double *)s->h;
    uint32_t *out = (uint32_t *)(s->h + 64);
    std::memcpy(slot, a, 3 * sizeof(double));
    if (b) std::memcpy(slot + 3, b, 3 * sizeof(double));
    const LaunchCfg cfg = launch_cfg(*ep);
    cudaError_t e;
    if (!b)
        e = launch_states(cfg, ep->view, ep->bp, slot, 1, kModeBool, can_pass_gate, 0.0, out, s->stream);
    else
        e = launch_segments(cfg, ep->view, ep->bp, slot, slot + 3, 1, can_pass_gate, out, s->stream);
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    EPP_CUDA(cudaStreamSynchronize(s->stream));
    *valid = (int)(*out & 1u);
    return EPP_OK;
}

} // namespace

// ---- misc -----------------------------------------------------------------------------------

extern "C" const char *epp_last_error(void) { return tl_error.c_str(); }

extern "C" const char *epp_version(void) { return "epp_b200 0.1 (sm_100a)"; }

extern "C" int epp_device_count(int *count)
{
    if (!count) return fail(EPP_ERR_INVALID_ARG);
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
        *count = 0;
        return cuda_fail(e, "cudaGetDeviceCount");
    }
    *count = n;
    return EPP_OK;
}

extern "C" uint64_t epp_kernel_launch_count(void) { return g_launches.load(); }

extern "C" int epp_set_tuning(const char *key, int value)
{
    if (!key) return fail(EPP_ERR_INVALID_ARG);
    std::atomic<int> *k = knob(key);
    if (!k) return fail(EPP_ERR_INVALID_ARG, std::string("unknown tuning key ") + key);
    k->store(value);
    return EPP_OK;
}

extern "C" int epp_get_tuning(const char *key, int *value)
{
    if (!key || !value) return fail(EPP_ERR_INVALID_ARG);
    std::atomic<int> *k = knob(key);
    if (!k) return fail(EPP_ERR_INVALID_ARG, std::string("unknown tuning key ") + key);
    *value = k->load();
    return EPP_OK;
}

// ---- world lifecycle ----------------------------------------------------------------------------

extern "C" int epp_world_create(double inflate_gate, double inflate_obstacle, epp_world **out)
{
    if (!out) return fail(EPP_ERR_INVALID_ARG);
    *out = new (std::nothrow) epp_world(inflate_gate, inflate_obstacle);
    return *out ? EPP_OK : fail(EPP_ERR_INVALID_ARG, "out of memory");
}

extern "C" int epp_world_destroy(epp_world *w)
{
    delete w;
    return EPP_OK;
}

#define EPP_MUTATE(expr)                              \
    if (!w) return fail(EPP_ERR_INVALID_ARG, "null world"); \
    std::lock_guard<std::mutex> lk(w->mu);            \
    int rc = (expr);                                  \
    if (rc == EPP_OK) w->dirty = true;                \
    return rc == EPP_OK ? EPP_OK : fail(rc)

extern "C" int epp_world_set_gate_geometry(epp_world *w, int type_id, const epp_obb_desc *d, int n)
{
    EPP_MUTATE(w->host.set_gate_geometry(type_id, d, n));
}

extern "C" int epp_world_set_obstacle_geometry(epp_world *w, const epp_obb_desc *d, int n)
{
    EPP_MUTATE(w->host.set_obstacle_geometry(d, n));
}

extern "C" int epp_world_reset(epp_world *w) { EPP_MUTATE((w->host.reset(), EPP_OK)); }

extern "C" int epp_world_add_gate(epp_world *w, int gate_id, const double pose7[7])
{
    if (!pose7) return fail(EPP_ERR_INVALID_ARG);
    EPP_MUTATE(w->host.add_gate(gate_id, pose7, false));
}

extern "C" int epp_world_update_gate(epp_world *w, int gate_id, const double pose7[7])
{
    if (!pose7) return fail(EPP_ERR_INVALID_ARG);
    EPP_MUTATE(w->host.add_gate(gate_id, pose7, true));
}

extern "C" int epp_world_add_obstacle(epp_world *w, int obstacle_id, const double pose6[6])
{
    if (!pose6) return fail(EPP_ERR_INVALID_ARG);
    EPP_MUTATE(w->host.add_obstacle(obstacle_id, pose6));
}

extern "C" int epp_world_set_broadphase(epp_world *w, epp_broadphase mode)
{
    if (mode < EPP_BP_AUTO || mode > EPP_BP_GRID) return fail(EPP_ERR_INVALID_ARG, "broadphase");
    EPP_MUTATE((w->bp_request = mode, EPP_OK));
}

extern "C" int epp_world_size(const epp_world *w, int *n)
{
    if (!w || !n) return fail(EPP_ERR_INVALID_ARG);
    std::lock_guard<std::mutex> lk(w->mu);
    *n = w->host.size();
    return EPP_OK;
}

extern "C" int epp_world_get_obb(const epp_world *w, int index, epp_obb *out)
{
    if (!w || !out) return fail(EPP_ERR_INVALID_ARG);
    std::lock_guard<std::mutex> lk(w->mu);
    int rc = w->host.get(index, out);
    return rc == EPP_OK ? EPP_OK : fail(rc);
}

extern "C" int epp_world_epoch(const epp_world *w, uint64_t *epoch)
{
    if (!w || !epoch) return fail(EPP_ERR_INVALID_ARG);
    std::lock_guard<std::mutex> lk(w->mu);
    *epoch = w->epoch;
    return EPP_OK;
}

extern "C" int epp_world_active_broadphase(const epp_world *w, int *mode)
{
    if (!w || !mode) return fail(EPP_ERR_INVALID_ARG);
    std::lock_guard<std::mutex> lk(w->mu);
    if (!w->current) return fail(EPP_ERR_NOT_COMMITTED);
    *mode = w->current->bp;
    return EPP_OK;
}

extern "C" int epp_world_commit(epp_world *w)
{
    if (!w) return fail(EPP_ERR_INVALID_ARG, "null world");
    // the epoch this commit retires is released AFTER the world's lock (declared first, destroyed last): freeing its
    // blob synchronises the device and must not stall checks that only want to take a snapshot
    std::shared_ptr<DeviceEpoch> retired;
    std::lock_guard<std::mutex> lk(w->mu);

    int device = 0;
    EPP_CUDA(cudaGetDevice(&device));
    std::vector<AabbRec> aabbs;
    std::vector<ObbRec> obbs;
    int rc = w->host.flatten(aabbs, obbs);
    if (rc) return fail(rc);
    const int M = (int)aabbs.size();
    if (M > 32 && knob_value("morton_sort")) morton_sort(aabbs, obbs); // device order only; the host registry keeps its own

    int bp = w->bp_request;
    // A bits row is summarised by ONE 32-bit word (bit w = word w of the row is non-zero) and the ray walk keeps its
    // non-zero words in a 32-bit set: rows of more than 32 words (1024 OBBs) cannot be represented.
    const int bits_max = std::min(knob_value("bits_max_obbs"), kBitsMaxObbs);
    if (bp == EPP_BP_AUTO) bp = (M <= bits_max) ? EPP_BP_BITS : EPP_BP_GRID;
    if (bp == EPP_BP_BITS && M > kBitsMaxObbs)
        return fail(EPP_ERR_INVALID_ARG, "EPP_BP_BITS supports at most 1024 OBBs (use EPP_BP_GRID or EPP_BP_AUTO)");

    BitsTables bits;
    GridTables grid;
    if (bp == EPP_BP_BITS) {
        int nb = knob_value("bits_bins");
        const int W = std::max(1, (M + 31) / 32);
        if (nb <= 0) nb = W == 1 ? 256 : (W <= 8 ? 128 : (W <= 32 ? 64 : 32));
        build_bits(aabbs, nb, bits, knob_value("bits_range_d"));
    }
    if (bp == EPP_BP_GRID)
        build_grid(aabbs, obbs, knob_value("grid_max_cells"), grid, (long)knob_value("grid_occ_kb") * 1024);

    // K2's per-call record fold (kernels.cu fold_ray_records) depends on the call only through canPassGate: both
    // variants are built here once per epoch so the kernel can take them with one bulk copy.
    std::vector<double> ray_recs;
    {
        bool zr = M > 0 && M <= 256;
        for (const ObbRec &o : obbs)
            if (o.rot[2] != 0.0 || o.rot[5] != 0.0 || o.rot[6] != 0.0 || o.rot[7] != 0.0 || o.rot[8] != 1.0) zr = false;
        if (zr && knob_value("segments_recs_at_commit")) {
            ray_recs.assign((size_t)2 * M * 21 + 2, 0.0);
            for (int cpg = 0; cpg < 2; ++cpg)
                for (int k = 0; k < M; ++k) {
                    double *r = ray_recs.data() + ((size_t)cpg * M + k) * 21;
                    const bool filling = (obbs[k].flags & kFlagFilling) != 0;
                    for (int d = 0; d < 3; ++d) {
                        r[d] = aabbs[k].mn[d];
                        r[3 + d] = aabbs[k].mx[d];
                        r[6 + d] = obbs[k].center[d];
                        const double h = obbs[k].half[d];
                        const double hi = h + obbs[k].inflate;
                        r[13 + d] = filling ? h : hi;
                        r[16 + d] = hi;
                    }
                    if (filling && cpg) r[0] = std::numeric_limits<double>::infinity();
                    r[9] = obbs[k].rot[0];
                    r[10] = obbs[k].rot[1];
                    r[11] = obbs[k].rot[3];
                    r[12] = obbs[k].rot[4];
                }
        }
    }

    // One blob: [aabb][obb][bits inside|min_le|max_ge][cell_start][cell_items], 256 B aligned parts.
    struct Part {
        const void *src;
        size_t bytes;
        size_t off;
    };
    std::vector<Part> parts = {
        {aabbs.data(), aabbs.size() * sizeof(AabbRec), 0},      {obbs.data(), obbs.size() * sizeof(ObbRec), 0},
        {bits.inside.data(), bits.inside.size() * 4, 0},        {bits.min_le.data(), bits.min_le.size() * 4, 0},
        {bits.max_ge.data(), bits.max_ge.size() * 4, 0},        {grid.cell_start.data(), grid.cell_start.size() * 4, 0},
        {grid.cell_items.data(), grid.cell_items.size() * 4, 0}, {grid.occ.data(), grid.occ.size() * 4, 0},
        {nullptr, 0, 0}, // (slot kept: the parts below are addressed by index)
        {grid.cell_recs.data(), grid.cell_recs.size() * 8, 0},
        {bits.range.data(), bits.range.size() * 4, 0},
        {ray_recs.data(), ray_recs.size() * 8, 0}};
    size_t total = 0;
    for (Part &p : parts) {
        p.off = total;
        total += align_up(std::max<size_t>(p.bytes, 16), 256);
    }
    std::shared_ptr<DeviceEpoch> ep(new DeviceEpoch);
    ep->device = device;
    EPP_CUDA(cudaMalloc(&ep->blob, total));
    EPP_CUDA(cudaMemset(ep->blob, 0, total));
    for (const Part &p : parts)
        if (p.bytes) EPP_CUDA(cudaMemcpy((unsigned char *)ep->blob + p.off, p.src, p.bytes, cudaMemcpyHostToDevice));
    EPP_CUDA(cudaDeviceGetAttribute(&ep->sm_count, cudaDevAttrMultiProcessorCount, device));

    unsigned char *base = (unsigned char *)ep->blob;
    WorldView &v = ep->view;
    v.aabb = (const AabbRec *)(base + parts[0].off);
    v.obb = (const ObbRec *)(base + parts[1].off);
    v.M = M;
    v.bits_inside = (const uint32_t *)(base + parts[2].off);
    v.bits_min_le = (const uint32_t *)(base + parts[3].off);
    v.bits_max_ge = (const uint32_t *)(base + parts[4].off);
    v.ray_recs = ray_recs.empty() ? nullptr : (const double *)(base + parts[11].off);
    v.bits_range = (const uint32_t *)(base + parts[10].off);
    v.range_d = bits.range.empty() ? 0 : bits.range_d;
    v.nb = bits.nb;
    v.W = bits.W;
    v.Ws = bits.Ws;
    v.kb = bits.bins;
    v.cmask_off = bits.cmask_off;
    v.cmask_words = knob_value("bits_cmask") ? bits.cmask_words : 0; // 0: per-axis summaries / full row scans (round-1 form)
    v.cn0 = bits.cn[0];
    v.cn1 = bits.cn[1];
    for (int a = 0; a < 3; ++a) v.cshift[a] = bits.cshift[a];
    v.zrot = 1;
    for (const ObbRec &o : obbs)
        if (o.rot[2] != 0.0 || o.rot[5] != 0.0 || o.rot[6] != 0.0 || o.rot[7] != 0.0 || o.rot[8] != 1.0) v.zrot = 0;
    v.cell_start = (const uint32_t *)(base + parts[5].off);
    v.cell_items = (const uint32_t *)(base + parts[6].off);
    v.gbins = grid.bins;
    v.cell_recs = (grid.cell_recs.empty() || !knob_value("grid_cell_recs")) ? nullptr : (const double *)(base + parts[9].off);
    v.occ = (const uint32_t *)(base + parts[7].off);
    v.occ_bytes = (uint32_t)align_up(grid.occ.size() * 4, 16); // parts are padded to 256 B
    v.occ_n0 = grid.occ_n[0];
    v.occ_n1 = grid.occ_n[1];
    for (int a = 0; a < 3; ++a) v.occ_shift[a] = grid.occ_shift[a];
    ep->bp = bp;
    ep->epoch = ++w->epoch;
    retired = std::move(w->current);
    w->current = ep; // the previous epoch is released when its last in-flight user drops it
    w->dirty = false;
    return EPP_OK;
}

extern "C" int epp_world_parse_gates_and_obstacles(epp_world *w, const double *gates7, int n_gates,
                                                   const double *obstacles6, int n_obstacles)
{
    if (!w || n_gates < 0 || n_obstacles < 0 || (n_gates && !gates7) || (n_obstacles && !obstacles6))
        return fail(EPP_ERR_INVALID_ARG);
    {
        std::lock_guard<std::mutex> lk(w->mu);
        w->host.reset(); // src/PathPlanner.cpp:63
        w->dirty = true;
        for (int i = 0; i < n_gates; ++i) {
            double g[7];
            std::memcpy(g, gates7 + 7 * i, sizeof g);
            g[2] = 0.0; // src/PathPlanner.cpp:68
            int rc = w->host.add_gate(i, g, false);
            if (rc) return fail(rc);
        }
        for (int i = 0; i < n_obstacles; ++i) {
            int rc = w->host.add_obstacle(i, obstacles6 + 6 * i);
            if (rc) return fail(rc);
        }
    }
    return epp_world_commit(w);
}

// ---- host-buffer checks ----------------------------------------------------------------------------

extern "C" int epp_check_states(epp_world *w, const double *xyz, size_t n, int can_pass_gate, uint32_t *mask)
{
    return check_host(w, xyz, nullptr, n, 0, kModeBool, can_pass_gate, 0.0, mask);
}

extern "C" int epp_check_states_mindist(epp_world *w, const double *xyz, size_t n, double min_distance,
                                        uint32_t *mask)
{
    return check_host(w, xyz, nullptr, n, 0, kModeMinDist, 0, min_distance, mask);
}

extern "C" int epp_check_segments(epp_world *w, const double *a, const double *b, size_t n, int can_pass_gate,
                                  uint32_t *mask)
{
    return check_host(w, a, b, n, 1, kModeBool, can_pass_gate, 0.0, mask);
}

// Mean wall time of n back-to-back epp_is_valid calls (microseconds per call), measured inside the library so that no
// binding overhead is in it: the number an OMPL planner thread sees per StateValidator::isValid.
extern "C" int epp_measure_single_latency(epp_world *w, int n_calls, double *us_per_call)
{
    if (!us_per_call || n_calls <= 0) return fail(EPP_ERR_INVALID_ARG);
    int valid = 0;
    double p[3] = {0.0, 0.0, 1.0};
    for (int i = 0; i < 16; ++i) { // warm-up: staging, first launch of the resident warp
        int rc = check_single(w, p, nullptr, 0, &valid);
        if (rc) return rc;
    }
    const auto t0 = Clock::now();
    for (int i = 0; i < n_calls; ++i) {
        p[0] = 0.01 * (double)(i & 127);
        int rc = check_single(w, p, nullptr, 0, &valid);
        if (rc) return rc;
    }
    *us_per_call = std::chrono::duration<double, std::micro>(Clock::now() - t0).count() / n_calls;
    return EPP_OK;
}

extern "C" int epp_is_valid(epp_world *w, const double p[3], int can_pass_gate, int *valid)
{
    return check_single(w, p, nullptr, can_pass_gate, valid);
}

extern "C" int epp_check_motion(epp_world *w, const double a[3], const double b[3], int can_pass_gate, int *valid)
{
    if (!b) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    return check_single(w, a, b, can_pass_gate, valid);
}

extern "C" int epp_check_trajectories(epp_world *w, const double *col_major, size_t n_traj, size_t rows, size_t ld,
                                      size_t traj_stride, int cols, double min_distance, uint32_t *traj_valid,
                                      uint32_t *sample_mask, int64_t *first_invalid)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n_traj == 0) return EPP_OK;
    if (!col_major || !traj_valid || cols < 7 || ld < rows || (n_traj > 1 && traj_stride < ld * (size_t)cols))
        return fail(EPP_ERR_INVALID_ARG, "trajectory matrix shape");
    DeviceGuard guard;
    EPP_CUDA(guard.enter(ep->device));
    Staging *s;
    rc = get_staging(ep->device, &s);
    if (rc) return rc;

    // Only the columns up to 6 of every matrix are needed on the device.
    const size_t span = (n_traj - 1) * traj_stride + 6 * ld + rows;
    const size_t in_bytes = span * sizeof(double);
    const size_t wpt = mask_words(rows);
    const size_t off_valid = align_up(in_bytes, 256);
    const size_t off_first = off_valid + align_up(mask_words(n_traj) * 4, 256);
    const size_t off_scratch = off_first + align_up(n_traj * 8, 256);
    const size_t off_smask = off_scratch + align_up(n_traj * 8, 256);
    const size_t smask_bytes = sample_mask ? n_traj * wpt * 4 : 0;
    rc = reserve_device(s, off_smask + smask_bytes);
    if (rc) return rc;
    const size_t out_bytes = mask_words(n_traj) * 4 + n_traj * 8 + smask_bytes;
    rc = reserve_host(s, 2 * kChunk + out_bytes + 512);
    if (rc) return rc;
    rc = upload(s, s->d, col_major, in_bytes, 0);
    if (rc) return rc;
    cudaError_t e = launch_trajectories(launch_cfg(*ep), ep->view, ep->bp, (const double *)s->d, n_traj, rows, ld,
                                        traj_stride, min_distance, (uint32_t *)(s->d + off_valid),
                                        sample_mask ? (uint32_t *)(s->d + off_smask) : nullptr,
                                        (int64_t *)(s->d + off_first), (long long *)(s->d + off_scratch), s->stream);
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    unsigned char *hb = s->h + 2 * kChunk;
    const size_t h_first = align_up(mask_words(n_traj) * 4, 8);
    const size_t h_smask = h_first + n_traj * 8;
    EPP_CUDA(cudaMemcpyAsync(hb, s->d + off_valid, mask_words(n_traj) * 4, cudaMemcpyDeviceToHost, s->stream));
    if (first_invalid)
        EPP_CUDA(cudaMemcpyAsync(hb + h_first, s->d + off_first, n_traj * 8, cudaMemcpyDeviceToHost, s->stream));
    if (sample_mask)
        EPP_CUDA(cudaMemcpyAsync(hb + h_smask, s->d + off_smask, smask_bytes, cudaMemcpyDeviceToHost, s->stream));
    EPP_CUDA(cudaStreamSynchronize(s->stream));
    std::memcpy(traj_valid, hb, mask_words(n_traj) * 4);
    if (first_invalid) std::memcpy(first_invalid, hb + h_first, n_traj * 8);
    if (sample_mask) std::memcpy(sample_mask, hb + h_smask, smask_bytes);
    return EPP_OK;
}

extern "C" int epp_check_poly_trajectories(epp_world *w, const double *coeffs, const double *durations,
                                           const double *dt, size_t n_traj, int n_segments, int n_coeffs,
                                           double min_distance, size_t max_samples, uint32_t *traj_valid,
                                           uint32_t *sample_mask, uint32_t *n_samples, int64_t *first_invalid)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n_traj == 0) return EPP_OK;
    if (!coeffs || !durations || !dt || !traj_valid || n_segments <= 0 || n_segments > 65535 || n_coeffs <= 0 ||
        max_samples == 0 || max_samples % 32)
        return fail(EPP_ERR_INVALID_ARG, "polynomial trajectory batch shape");
    DeviceGuard guard;
    EPP_CUDA(guard.enter(ep->device));
    Staging *s;
    rc = get_staging(ep->device, &s);
    if (rc) return rc;

    const size_t c_bytes = n_traj * (size_t)n_segments * 3 * (size_t)n_coeffs * 8;
    const size_t d_bytes = n_traj * (size_t)n_segments * 8;
    const size_t t_bytes = n_traj * 8;
    const size_t wpt = max_samples / 32;
    const size_t off_d = align_up(c_bytes, 256);
    const size_t off_t = off_d + align_up(d_bytes, 256);
    const size_t off_valid = off_t + align_up(t_bytes, 256);
    const size_t off_first = off_valid + align_up(mask_words(n_traj) * 4, 256);
    const size_t off_scratch = off_first + align_up(n_traj * 8, 256);
    const size_t off_ns = off_scratch + align_up(n_traj * 8, 256);
    const size_t off_smask = off_ns + align_up(n_traj * 4, 256);
    const size_t smask_bytes = sample_mask ? n_traj * wpt * 4 : 0;
    rc = reserve_device(s, off_smask + smask_bytes);
    if (rc) return rc;
    const size_t out_bytes = mask_words(n_traj) * 4 + n_traj * 12 + smask_bytes;
    rc = reserve_host(s, 2 * kChunk + out_bytes + 512);
    if (rc) return rc;
    if ((rc = upload(s, s->d, coeffs, c_bytes, 0))) return rc;
    if ((rc = upload(s, s->d + off_d, durations, d_bytes, 0))) return rc;
    if ((rc = upload(s, s->d + off_t, dt, t_bytes, 0))) return rc;
    cudaError_t e = launch_poly_trajectories(
        launch_cfg(*ep), ep->view, ep->bp, (const double *)s->d, (const double *)(s->d + off_d),
        (const double *)(s->d + off_t), n_traj, n_segments, n_coeffs, min_distance, max_samples,
        (uint32_t *)(s->d + off_valid), sample_mask ? (uint32_t *)(s->d + off_smask) : nullptr,
        (uint32_t *)(s->d + off_ns), (int64_t *)(s->d + off_first), (long long *)(s->d + off_scratch), s->stream);
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    unsigned char *hb = s->h + 2 * kChunk;
    const size_t h_first = align_up(mask_words(n_traj) * 4, 8);
    const size_t h_ns = h_first + n_traj * 8;
    const size_t h_smask = h_ns + align_up(n_traj * 4, 8);
    EPP_CUDA(cudaMemcpyAsync(hb, s->d + off_valid, mask_words(n_traj) * 4, cudaMemcpyDeviceToHost, s->stream));
    EPP_CUDA(cudaMemcpyAsync(hb + h_first, s->d + off_first, n_traj * 8, cudaMemcpyDeviceToHost, s->stream));
    EPP_CUDA(cudaMemcpyAsync(hb + h_ns, s->d + off_ns, n_traj * 4, cudaMemcpyDeviceToHost, s->stream));
    if (sample_mask)
        EPP_CUDA(cudaMemcpyAsync(hb + h_smask, s->d + off_smask, smask_bytes, cudaMemcpyDeviceToHost, s->stream));
    EPP_CUDA(cudaStreamSynchronize(s->stream));
    std::memcpy(traj_valid, hb, mask_words(n_traj) * 4);
    if (first_invalid) std::memcpy(first_invalid, hb + h_first, n_traj * 8);
    if (n_samples) std::memcpy(n_samples, hb + h_ns, n_traj * 4);
    if (sample_mask) std::memcpy(sample_mask, hb + h_smask, smask_bytes);
    return EPP_OK;
}

// ---- device-buffer checks (asynchronous, capture-safe: launches only) ---------------------------------

extern "C" int epp_check_states_dev(epp_world *w, const double *d_xyz, size_t n, int can_pass_gate, uint32_t *d_mask,
                                    void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n && (!d_xyz || !d_mask)) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    cudaError_t e = launch_states(launch_cfg(*ep), ep->view, ep->bp, d_xyz, n, kModeBool, can_pass_gate, 0.0, d_mask,
                                  (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

// Several INDEPENDENT batches in one call (the planner's sample batches): launched back to back on `stream`; from the
// second on, a batch does not wait for the previous one before it starts streaming its states (programmatic dependent
// launch with the wait at the kernel's end).  The caller guarantees what makes that safe: every input buffer is
// complete before the call, and no buffer of one batch overlaps a mask of another.
extern "C" int epp_check_states_multi_dev(epp_world *w, const double *const *d_xyz, const size_t *n, uint32_t *const *d_mask,
                                          int batches, int can_pass_gate, void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (batches < 0 || (batches && (!d_xyz || !n || !d_mask))) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    for (int i = 0; i < batches; ++i) {
        if (n[i] && (!d_xyz[i] || !d_mask[i])) return fail(EPP_ERR_INVALID_ARG, "null buffer");
        for (int j = 0; j < i; ++j) { // masks must not overlap each other or any input
            const char *mi = (const char *)d_mask[i], *mj = (const char *)d_mask[j];
            const size_t bi = (n[i] + 31) / 32 * 4, bj = (n[j] + 31) / 32 * 4;
            if (mi < mj + bj && mj < mi + bi) return fail(EPP_ERR_INVALID_ARG, "masks of two batches overlap");
        }
    }
    LaunchCfg cfg = launch_cfg(*ep);
    const int want_late = knob_value("states_late_wait");
    cfg.states_multi = (want_late && batches > 1) ? std::max(1, std::min(8, knob_value("states_multi_blocks_per_sm"))) : 0;
    cfg.states_multi_grid = knob_value("states_multi_grid_per_sm");
    for (int i = 0; i < batches; ++i) {
        cfg.states_late_wait = (i > 0 && want_late && n[i - 1] > 0) ? 1 : 0; // only behind a launch of this very call
        cudaError_t e = launch_states(cfg, ep->view, ep->bp, d_xyz[i], n[i], kModeBool, can_pass_gate, 0.0, d_mask[i],
                                      (cudaStream_t)stream);
        if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    }
    return EPP_OK;
}

extern "C" int epp_check_states_mindist_dev(epp_world *w, const double *d_xyz, size_t n, double min_distance,
                                            uint32_t *d_mask, void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n && (!d_xyz || !d_mask)) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    cudaError_t e = launch_states(launch_cfg(*ep), ep->view, ep->bp, d_xyz, n, kModeMinDist, 0, min_distance, d_mask,
                                  (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_check_segments_dev(epp_world *w, const double *d_a, const double *d_b, size_t n, int can_pass_gate,
                                      uint32_t *d_mask, void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n && (!d_a || !d_b || !d_mask)) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    cudaError_t e = launch_segments(launch_cfg(*ep), ep->view, ep->bp, d_a, d_b, n, can_pass_gate, d_mask,
                                    (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_check_segments_multi_dev(epp_world *w, const double *const *d_a, const double *const *d_b,
                                            const size_t *n, uint32_t *const *d_mask, int batches, int can_pass_gate,
                                            void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (batches < 0 || (batches && (!d_a || !d_b || !n || !d_mask))) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    for (int i = 0; i < batches; ++i) {
        if (n[i] && (!d_a[i] || !d_b[i] || !d_mask[i])) return fail(EPP_ERR_INVALID_ARG, "null buffer");
        for (int j = 0; j < i; ++j) { // masks must not overlap each other
            const char *mi = (const char *)d_mask[i], *mj = (const char *)d_mask[j];
            const size_t bi = (n[i] + 31) / 32 * 4, bj = (n[j] + 31) / 32 * 4;
            if (mi < mj + bj && mj < mi + bi) return fail(EPP_ERR_INVALID_ARG, "masks of two batches overlap");
        }
    }
    LaunchCfg cfg = launch_cfg(*ep);
    const int want_late = knob_value("states_late_wait");
    cfg.segments_multi = (want_late && batches > 1) ? std::max(1, std::min(4, knob_value("segments_multi_blocks_per_sm"))) : 0;
    for (int i = 0; i < batches; ++i) {
        cfg.segments_late_wait = (i > 0 && want_late && n[i - 1] > 0) ? 1 : 0; // only behind a launch of this very call
        cudaError_t e = launch_segments(cfg, ep->view, ep->bp, d_a[i], d_b[i], n[i], can_pass_gate, d_mask[i],
                                        (cudaStream_t)stream);
        if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    }
    return EPP_OK;
}

extern "C" int epp_check_trajectories_dev(epp_world *w, const double *d_col_major, size_t n_traj, size_t rows,
                                          size_t ld, size_t traj_stride, int cols, double min_distance,
                                          uint32_t *d_traj_valid, uint32_t *d_sample_mask, int64_t *d_first_invalid,
                                          void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n_traj == 0) return EPP_OK;
    if (!d_col_major || !d_traj_valid || !d_first_invalid || cols < 7 || ld < rows)
        return fail(EPP_ERR_INVALID_ARG, "trajectory matrix shape (d_first_invalid is required on the device path)");
    // d_first_invalid doubles as the atomicMin scratch (same width), finalised in place.
    cudaError_t e = launch_trajectories(launch_cfg(*ep), ep->view, ep->bp, d_col_major, n_traj, rows, ld, traj_stride,
                                        min_distance, d_traj_valid, d_sample_mask, d_first_invalid,
                                        (long long *)d_first_invalid, (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_check_poly_trajectories_dev(epp_world *w, const double *d_coeffs, const double *d_durations,
                                               const double *d_dt, size_t n_traj, int n_segments, int n_coeffs,
                                               double min_distance, size_t max_samples, uint32_t *d_traj_valid,
                                               uint32_t *d_sample_mask, uint32_t *d_n_samples,
                                               int64_t *d_first_invalid, void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n_traj == 0) return EPP_OK;
    if (!d_coeffs || !d_durations || !d_dt || !d_traj_valid || !d_first_invalid || n_segments <= 0 ||
        n_segments > 65535 || n_coeffs <= 0 || max_samples == 0 || max_samples % 32)
        return fail(EPP_ERR_INVALID_ARG, "polynomial trajectory batch shape");
    cudaError_t e = launch_poly_trajectories(launch_cfg(*ep), ep->view, ep->bp, d_coeffs, d_durations, d_dt, n_traj,
                                             n_segments, n_coeffs, min_distance, max_samples, d_traj_valid,
                                             d_sample_mask, d_n_samples, d_first_invalid, (long long *)d_first_invalid,
                                             (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

// ---- mask exchange fused into the check kernels ----------------------------------------------------------

struct epp_gather {
    int rank = 0, world = 1, device = 0, n_buffers = 1;
    size_t slot_words = 0;
    // [control words: 256 B][n_buffers x world x slot_words words][staging: 2 x world x (1 + ceil(slot_words / 3)) quads]
    unsigned char *base = nullptr;
    unsigned char *peer_base[kMaxRanks] = {}; // every rank's allocation in this process's address space
    bool opened[kMaxRanks] = {};              // peer_base[p] came from cudaIpcOpenMemHandle
    bool connected = false;
    int last_buffer = -1;                     // buffer of the previous gathering call
    bool deferred = false;                    // deferred form (epp_gather_set_deferred)
    uint32_t *prev_out = nullptr;             // deferred form: the previous call's gathered buffer, still to be completed
    uint64_t timeout_ns = 2000000000ull;
    size_t buffer_bytes() const { return (size_t)world * slot_words * 4; }
    uint32_t *buffer_of(unsigned char *b, int i) const { return (uint32_t *)(b + 256 + (size_t)i * buffer_bytes()); }
    size_t staging_off() const { return (256 + (size_t)n_buffers * buffer_bytes() + 255) / 256 * 256; }
    size_t row_quads() const { return 2 + (slot_words + 2) / 3; }
    size_t staging_bytes() const { return (size_t)2 * world * row_quads() * 16; }
};

namespace {

constexpr size_t kGatherCtlOff = 0; // control words: blocks done, calls completed, time-outs

int gather_view(epp_gather *g, int buffer, size_t n, GatherView &gv)
{
    if (!g) return fail(EPP_ERR_INVALID_ARG, "null gather");
    if (!g->connected) return fail(EPP_ERR_INVALID_ARG, "gather is not connected to its peers");
    if (buffer < 0 || buffer >= g->n_buffers) return fail(EPP_ERR_INVALID_ARG, "gather buffer index");
    if ((n + 31) / 32 > g->slot_words) return fail(EPP_ERR_INVALID_ARG, "more units than the gather slot holds");
    for (int p = 0; p < kMaxRanks; ++p)
        gv.peer_ll[p] = p < g->world ? (uint4 *)(g->peer_base[p] + g->staging_off()) : nullptr;
    gv.my_ll = (const uint4 *)(g->base + g->staging_off());
    gv.row_quads = g->row_quads();
    for (int p = 0; p < kMaxRanks; ++p) gv.peer_out[p] = p < g->world ? g->buffer_of(g->peer_base[p], buffer) : nullptr;
    // Large slots are copied straight into the peers' gathered buffers (k_gather_bulk): a peer may be one call ahead,
    // so the buffer of the previous call must not be the target of this one.
    const int bulk_min = knob_value("gather_bulk_min_words");
    if (g->world > 1 && bulk_min > 0 && g->slot_words >= (size_t)bulk_min) {
        if (g->n_buffers < 2 || buffer == g->last_buffer)
            return fail(EPP_ERR_INVALID_ARG, "gather slots of this size are written by the peers directly: use n_buffers >= 2 "
                                             "and a different buffer than the previous call");
    }
    g->last_buffer = buffer;
    gv.out = g->buffer_of(g->base, buffer);
    const bool bulk = g->world > 1 && bulk_min > 0 && g->slot_words >= (size_t)bulk_min;
    gv.deferred = (g->deferred && g->world > 1 && !bulk) ? 1 : 0;
    gv.prev_out = gv.deferred ? g->prev_out : nullptr;
    if (gv.deferred) g->prev_out = gv.out;
    gv.ctl = (uint32_t *)(g->base + kGatherCtlOff);
    gv.slot_words = g->slot_words;
    gv.my_words = (n + 31) / 32;
    gv.timeout_ns = g->timeout_ns;
    gv.world = g->world;
    gv.rank = g->rank;
    gv.debug = knob_value("gather_debug");
    return EPP_OK;
}

} // namespace

extern "C" int epp_gather_create(int rank, int world_size, size_t slot_words, int n_buffers, epp_gather **out)
{
    if (!out || world_size < 1 || world_size > kMaxRanks || rank < 0 || rank >= world_size || n_buffers < 1 ||
        n_buffers > 64 || slot_words == 0 || slot_words % 4)
        return fail(EPP_ERR_INVALID_ARG, "gather shape (slot_words must be a positive multiple of 4, world_size <= 8)");
    std::unique_ptr<epp_gather> g(new (std::nothrow) epp_gather);
    if (!g) return fail(EPP_ERR_INVALID_ARG, "out of memory");
    g->rank = rank;
    g->world = world_size;
    g->slot_words = slot_words;
    g->n_buffers = n_buffers;
    EPP_CUDA(cudaGetDevice(&g->device));
    const size_t bytes = g->staging_off() + g->staging_bytes();
    EPP_CUDA(cudaMalloc((void **)&g->base, bytes));
    EPP_CUDA(cudaMemset(g->base, 0, bytes));
    EPP_CUDA(cudaDeviceSynchronize());
    g->peer_base[rank] = g->base;
    g->connected = world_size == 1;
    *out = g.release();
    return EPP_OK;
}

extern "C" int epp_gather_destroy(epp_gather *g)
{
    if (!g) return EPP_OK;
    int prev = -1;
    cudaGetDevice(&prev);
    cudaSetDevice(g->device);
    cudaDeviceSynchronize();
    for (int p = 0; p < g->world; ++p)
        if (g->opened[p]) cudaIpcCloseMemHandle(g->peer_base[p]);
    if (g->base) cudaFree(g->base);
    if (prev >= 0) cudaSetDevice(prev);
    delete g;
    return EPP_OK;
}

extern "C" int epp_gather_export(epp_gather *g, unsigned char handle[EPP_GATHER_HANDLE_BYTES])
{
    static_assert(sizeof(cudaIpcMemHandle_t) == EPP_GATHER_HANDLE_BYTES, "CUDA IPC handle size");
    if (!g || !handle) return fail(EPP_ERR_INVALID_ARG, "null gather");
    cudaIpcMemHandle_t h;
    EPP_CUDA(cudaIpcGetMemHandle(&h, g->base));
    std::memcpy(handle, &h, sizeof h);
    return EPP_OK;
}

extern "C" int epp_gather_connect(epp_gather *g, const unsigned char *handles)
{
    if (!g || !handles) return fail(EPP_ERR_INVALID_ARG, "null gather");
    DeviceGuard guard;
    EPP_CUDA(guard.enter(g->device));
    for (int p = 0; p < g->world; ++p) {
        if (p == g->rank || g->peer_base[p]) continue;
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handles + (size_t)p * EPP_GATHER_HANDLE_BYTES, sizeof h);
        void *ptr = nullptr;
        EPP_CUDA(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
        g->peer_base[p] = (unsigned char *)ptr;
        g->opened[p] = true;
    }
    g->connected = true;
    return EPP_OK;
}

extern "C" int epp_gather_connect_local(epp_gather *const *all, int world_size)
{
    if (!all || world_size < 1 || world_size > kMaxRanks) return fail(EPP_ERR_INVALID_ARG, "gather list");
    for (int r = 0; r < world_size; ++r)
        if (!all[r] || all[r]->rank != r || all[r]->world != world_size || all[r]->slot_words != all[0]->slot_words ||
            all[r]->n_buffers != all[0]->n_buffers)
            return fail(EPP_ERR_INVALID_ARG, "gather handles must be given in rank order and share one shape");
    int prev = -1;
    EPP_CUDA(cudaGetDevice(&prev));
    for (int r = 0; r < world_size; ++r) {
        EPP_CUDA(cudaSetDevice(all[r]->device));
        for (int p = 0; p < world_size; ++p) {
            if (all[p]->device != all[r]->device) {
                cudaError_t e = cudaDeviceEnablePeerAccess(all[p]->device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
                else if (e != cudaSuccess) {
                    cudaSetDevice(prev);
                    return cuda_fail(e, "cudaDeviceEnablePeerAccess");
                }
            }
            all[r]->peer_base[p] = all[p]->base;
        }
        all[r]->connected = true;
    }
    EPP_CUDA(cudaSetDevice(prev));
    return EPP_OK;
}

extern "C" int epp_gather_buffer(epp_gather *g, int buffer, uint32_t **d_words)
{
    if (!g || !d_words || buffer < 0 || buffer >= g->n_buffers) return fail(EPP_ERR_INVALID_ARG, "gather buffer index");
    *d_words = g->buffer_of(g->base, buffer);
    return EPP_OK;
}

extern "C" int epp_gather_set_timeout_ms(epp_gather *g, unsigned ms)
{
    if (!g) return fail(EPP_ERR_INVALID_ARG, "null gather");
    g->timeout_ns = (uint64_t)ms * 1000000ull;
    return EPP_OK;
}

extern "C" int epp_gather_set_deferred(epp_gather *g, int on)
{
    if (!g) return fail(EPP_ERR_INVALID_ARG, "null gather");
    if (g->prev_out) return fail(EPP_ERR_INVALID_ARG, "flush the gather before changing its mode");
    g->deferred = on != 0;
    return EPP_OK;
}

extern "C" int epp_gather_flush(epp_gather *g, void *stream)
{
    if (!g) return fail(EPP_ERR_INVALID_ARG, "null gather");
    if (!g->connected) return fail(EPP_ERR_INVALID_ARG, "gather is not connected to its peers");
    if (!g->deferred || !g->prev_out) return EPP_OK; // nothing outstanding
    GatherView gv;
    int rc = gather_view(g, g->last_buffer, 0, gv);
    if (rc) return rc;
    g->prev_out = nullptr; // gather_view has just set it to this buffer again
    int device = 0, sms = 148;
    EPP_CUDA(cudaGetDevice(&device));
    EPP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    LaunchCfg cfg;
    cfg.sm_count = sms;
    const int sm_limit = knob_value("sm_limit");
    if (sm_limit > 0 && sm_limit < cfg.sm_count) cfg.sm_count = sm_limit;
    cudaError_t e = launch_gather_deferred(cfg, gv, 1, (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_gather_status(epp_gather *g, uint64_t *calls, uint64_t *timeouts)
{
    if (!g) return fail(EPP_ERR_INVALID_ARG, "null gather");
    DeviceGuard guard;
    EPP_CUDA(guard.enter(g->device));
    uint32_t ctl[4];
    EPP_CUDA(cudaDeviceSynchronize());
    EPP_CUDA(cudaMemcpy(ctl, g->base + kGatherCtlOff, sizeof ctl, cudaMemcpyDeviceToHost));
    if (calls) *calls = ctl[1];
    if (timeouts) *timeouts = ctl[2];
    return EPP_OK;
}

extern "C" int epp_check_states_gather_dev(epp_world *w, const double *d_xyz, size_t n, int can_pass_gate,
                                           epp_gather *g, int buffer, void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n && !d_xyz) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    GatherView gv;
    rc = gather_view(g, buffer, n, gv);
    if (rc) return rc;
    cudaError_t e = launch_states(launch_cfg(*ep), ep->view, ep->bp, d_xyz, n, kModeBool, can_pass_gate, 0.0,
                                  gv.out + (size_t)gv.rank * gv.slot_words, (cudaStream_t)stream, g->world > 1 ? &gv : nullptr);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_check_states_mindist_gather_dev(epp_world *w, const double *d_xyz, size_t n, double min_distance,
                                                   epp_gather *g, int buffer, void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n && !d_xyz) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    GatherView gv;
    rc = gather_view(g, buffer, n, gv);
    if (rc) return rc;
    cudaError_t e = launch_states(launch_cfg(*ep), ep->view, ep->bp, d_xyz, n, kModeMinDist, 0, min_distance,
                                  gv.out + (size_t)gv.rank * gv.slot_words, (cudaStream_t)stream, g->world > 1 ? &gv : nullptr);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_check_segments_gather_dev(epp_world *w, const double *d_a, const double *d_b, size_t n,
                                             int can_pass_gate, epp_gather *g, int buffer, void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n && (!d_a || !d_b)) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    GatherView gv;
    rc = gather_view(g, buffer, n, gv);
    if (rc) return rc;
    cudaError_t e = launch_segments(launch_cfg(*ep), ep->view, ep->bp, d_a, d_b, n, can_pass_gate,
                                    gv.out + (size_t)gv.rank * gv.slot_words, (cudaStream_t)stream, g->world > 1 ? &gv : nullptr);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

// ---- min-snap trajectory generation ------------------------------------------------------------------------

extern "C" int epp_segment_times(const double *waypoints, int n_vertices, double v_max, double a_max, double *durations)
{
    if (!waypoints || !durations || n_vertices < 2 || !(v_max > 0.0) || !(a_max > 0.0))
        return fail(EPP_ERR_INVALID_ARG, "segment times");
    const double magic_fabian_constant = 6.5; // vertex.h default
    for (int i = 0; i + 1 < n_vertices; ++i) {
        const double *s = waypoints + 3 * i, *e = s + 3;
        const double dx = e[0] - s[0], dy = e[1] - s[1], dz = e[2] - s[2];
        const double distance = std::sqrt(dx * dx + dy * dy + dz * dz);
        durations[i] = distance / v_max * 2 * (1.0 + magic_fabian_constant * v_max / a_max * std::exp(-distance / v_max * 2));
    }
    return EPP_OK;
}

extern "C" int epp_minsnap_dev(const double *d_waypoints, const double *d_durations, const double *d_start_vel_acc,
                               size_t n_traj, int n_segments, double *d_coeffs, void *stream)
{
    if (n_traj && (!d_waypoints || !d_durations || !d_coeffs)) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    if (n_segments < 1 || n_segments > 128) return fail(EPP_ERR_INVALID_ARG, "1 to 128 segments per trajectory");
    cudaError_t e = launch_minsnap(d_waypoints, d_durations, d_start_vel_acc, n_traj, n_segments, d_coeffs,
                                   (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_minsnap(const double *waypoints, const double *durations, const double *start_vel_acc, size_t n_traj,
                           int n_segments, double *coeffs)
{
    if (n_traj == 0) return EPP_OK;
    if (!waypoints || !durations || !coeffs) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    if (n_segments < 1 || n_segments > 128) return fail(EPP_ERR_INVALID_ARG, "1 to 128 segments per trajectory");
    int device = 0;
    EPP_CUDA(cudaGetDevice(&device));
    Staging *s;
    int rc = get_staging(device, &s);
    if (rc) return rc;
    const size_t K = (size_t)n_segments;
    const size_t b_wp = n_traj * (K + 1) * 3 * 8, b_T = n_traj * K * 8, b_va = start_vel_acc ? n_traj * 6 * 8 : 0;
    const size_t b_c = n_traj * K * 30 * 8;
    const size_t o_T = align_up(b_wp, 256), o_va = o_T + align_up(b_T, 256), o_c = o_va + align_up(b_va, 256);
    rc = reserve_device(s, o_c + b_c);
    if (rc) return rc;
    rc = reserve_host(s, 2 * kChunk);
    if (rc) return rc;
    if ((rc = upload(s, s->d, waypoints, b_wp, 0))) return rc;
    if ((rc = upload(s, s->d + o_T, durations, b_T, 0))) return rc;
    if (b_va && (rc = upload(s, s->d + o_va, start_vel_acc, b_va, 0))) return rc;
    cudaError_t e = launch_minsnap((const double *)s->d, (const double *)(s->d + o_T),
                                   b_va ? (const double *)(s->d + o_va) : nullptr, n_traj, n_segments,
                                   (double *)(s->d + o_c), s->stream);
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    EPP_CUDA(cudaMemcpyAsync(coeffs, s->d + o_c, b_c, cudaMemcpyDeviceToHost, s->stream));
    EPP_CUDA(cudaStreamSynchronize(s->stream));
    return EPP_OK;
}

extern "C" int epp_sample_trajectories_dev(const double *d_coeffs, const double *d_durations, const double *d_dt,
                                           const double *d_time_offsets, size_t n_traj, int n_segments, int n_coeffs,
                                           size_t ld, double *d_out, uint32_t *d_n_rows, void *stream)
{
    if (n_traj == 0) return EPP_OK;
    if (!d_coeffs || !d_durations || !d_dt || !d_out || n_segments < 1 || n_coeffs < 1 || ld == 0)
        return fail(EPP_ERR_INVALID_ARG, "trajectory sampling arguments");
    cudaError_t e = launch_sample_poly(d_coeffs, d_durations, d_dt, d_time_offsets, n_traj, n_segments, n_coeffs, ld, d_out,
                                       d_n_rows, (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_sample_trajectories(const double *coeffs, const double *durations, const double *dt,
                                       const double *time_offsets, size_t n_traj, int n_segments, int n_coeffs, size_t ld,
                                       double *out, uint32_t *n_rows)
{
    if (n_traj == 0) return EPP_OK;
    if (!coeffs || !durations || !dt || !out || n_segments < 1 || n_coeffs < 1 || ld == 0)
        return fail(EPP_ERR_INVALID_ARG, "trajectory sampling arguments");
    int device = 0;
    EPP_CUDA(cudaGetDevice(&device));
    Staging *s;
    int rc = get_staging(device, &s);
    if (rc) return rc;
    const size_t K = (size_t)n_segments;
    const size_t b_c = n_traj * K * 3 * (size_t)n_coeffs * 8, b_T = n_traj * K * 8, b_dt = n_traj * 8;
    const size_t b_off = time_offsets ? n_traj * 8 : 0, b_out = n_traj * 10 * ld * 8, b_n = n_traj * 4;
    const size_t o_T = align_up(b_c, 256), o_dt = o_T + align_up(b_T, 256), o_off = o_dt + align_up(b_dt, 256);
    const size_t o_out = o_off + align_up(b_off, 256), o_n = o_out + align_up(b_out, 256);
    rc = reserve_device(s, o_n + b_n);
    if (rc) return rc;
    rc = reserve_host(s, 2 * kChunk + b_n);
    if (rc) return rc;
    if ((rc = upload(s, s->d, coeffs, b_c, 0))) return rc;
    if ((rc = upload(s, s->d + o_T, durations, b_T, 0))) return rc;
    if ((rc = upload(s, s->d + o_dt, dt, b_dt, 0))) return rc;
    if (b_off && (rc = upload(s, s->d + o_off, time_offsets, b_off, 0))) return rc;
    cudaError_t e = launch_sample_poly((const double *)s->d, (const double *)(s->d + o_T), (const double *)(s->d + o_dt),
                                       b_off ? (const double *)(s->d + o_off) : nullptr, n_traj, n_segments, n_coeffs, ld,
                                       (double *)(s->d + o_out), (uint32_t *)(s->d + o_n), s->stream);
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    EPP_CUDA(cudaMemcpyAsync(out, s->d + o_out, b_out, cudaMemcpyDeviceToHost, s->stream));
    uint32_t *h_n = (uint32_t *)(s->h + 2 * kChunk);
    EPP_CUDA(cudaMemcpyAsync(h_n, s->d + o_n, b_n, cudaMemcpyDeviceToHost, s->stream));
    EPP_CUDA(cudaStreamSynchronize(s->stream));
    if (n_rows) std::memcpy(n_rows, h_n, b_n);
    return EPP_OK;
}

// ---- multi-GPU helpers, generators ---------------------------------------------------------------------

extern "C" int epp_shard_range(size_t n, int rank, int world_size, size_t align, size_t *begin, size_t *end)
{
    if (!begin || !end || world_size <= 0 || rank < 0 || rank >= world_size || align == 0 || align % 32)
        return fail(EPP_ERR_INVALID_ARG, "shard range");
    const size_t blocks = (n + align - 1) / align;
    const size_t per = (blocks + (size_t)world_size - 1) / (size_t)world_size * align;
    *begin = std::min(n, per * (size_t)rank);
    *end = std::min(n, *begin + per);
    return EPP_OK;
}

extern "C" int epp_mask_summary_dev(const uint32_t *d_mask, size_t n_words, uint64_t first_word,
                                    uint64_t *d_popcount_xor, void *stream)
{
    if (!d_popcount_xor || (n_words && !d_mask)) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    LaunchCfg cfg;
    int device = 0;
    EPP_CUDA(cudaGetDevice(&device));
    EPP_CUDA(cudaDeviceGetAttribute(&cfg.sm_count, cudaDevAttrMultiProcessorCount, device));
    cudaError_t e = launch_mask_summary(cfg, d_mask, n_words, first_word, d_popcount_xor, (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_count_candidates_dev(epp_world *w, const double *d_a, const double *d_b, size_t n,
                                       uint64_t *d_visited_hits, void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (!d_visited_hits || (n && !d_a)) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    DeviceGuard guard;
    EPP_CUDA(guard.enter(ep->device));
    cudaError_t e = launch_count_candidates(launch_cfg(*ep), ep->view, ep->bp, d_a, d_b, n, d_visited_hits, (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_gen_states_dev(uint64_t seed, uint64_t first, size_t n, const double lo[3], const double hi[3],
                                  double *d_xyz, void *stream)
{
    if (!lo || !hi || (n && !d_xyz)) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    cudaError_t e = launch_gen_states(seed, first, n, lo, hi, d_xyz, (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

extern "C" int epp_gen_segments_dev(uint64_t seed, uint64_t first, size_t n, const double lo[3], const double hi[3],
                                    double max_len, double *d_a, double *d_b, void *stream)
{
    if (!lo || !hi || (n && (!d_a || !d_b))) return fail(EPP_ERR_INVALID_ARG, "null buffer");
    cudaError_t e = launch_gen_segments(seed, first, n, lo, hi, max_len, d_a, d_b, (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}

// ---- allocation-free variant of the fused polynomial sweep ------------------------------------------------
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// K3b keeps one "first invalid row" word per trajectory that its blocks lower with atomicMin and a last kernel turns
// into the validity bits.  epp_check_poly_trajectories_dev takes that word array from the caller (d_first_invalid is
// required there); here it lives in a file-scope device array of EPP_STATIC_MAX_TRAJECTORIES words (512 KiB, held by
// every GPU that loads the library until the process exits), so the caller passes inputs and outputs only and
// d_first_invalid becomes optional.  The array is shared by every call and keeps the last call's contents: the launch
// sequence refills the words it uses with LLONG_MAX first (k_fill_i64), and calls that overlap on the GPU (two
// streams, two threads) race on it — serialise them or use the caller-scratch entry point.
//
// g_static_first_row is a host-side stand-in in this file's host code, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, asked once per device with that device current.
static __device__ __align__(256) long long g_static_first_row[EPP_STATIC_MAX_TRAJECTORIES];

namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
long long *g_static_ptr[kStaticMaxDevices] = {};

// Device address of g_static_first_row on the CURRENT device (which the caller has made `device`).
int static_first_row(int device, long long **out)
{
    if (device < 0 || device >= kStaticMaxDevices) return fail(EPP_ERR_INVALID_ARG, "device ordinal beyond 63");
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_static_first_row);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_first_row)");
        if (!p) return fail(EPP_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        g_static_ptr[device] = static_cast<long long *>(p);
    }
    *out = g_static_ptr[device];
    return EPP_OK;
}
} // namespace

extern "C" int epp_check_poly_trajectories_static_dev(epp_world *w, const double *d_coeffs, const double *d_durations,
                                                      const double *d_dt, size_t n_traj, int n_segments, int n_coeffs,
                                                      double min_distance, size_t max_samples, uint32_t *d_traj_valid,
                                                      uint32_t *d_sample_mask, uint32_t *d_n_samples,
                                                      int64_t *d_first_invalid, void *stream)
{
    std::shared_ptr<DeviceEpoch> ep;
    int rc = snapshot(w, ep);
    if (rc) return rc;
    if (n_traj > (size_t)EPP_STATIC_MAX_TRAJECTORIES)
        return fail(EPP_ERR_INVALID_ARG, "n_traj = " + std::to_string(n_traj) + " exceeds EPP_STATIC_MAX_TRAJECTORIES = " +
                                             std::to_string(EPP_STATIC_MAX_TRAJECTORIES) +
                                             " (the static scratch array); split the batch or use epp_check_poly_trajectories_dev");
    if (n_traj == 0) return EPP_OK;
    if (!d_coeffs || !d_durations || !d_dt || !d_traj_valid || n_segments <= 0 || n_segments > 65535 || n_coeffs <= 0 ||
        max_samples == 0 || max_samples % 32)
        return fail(EPP_ERR_INVALID_ARG, "polynomial trajectory batch shape");
    int device = -1;
    EPP_CUDA(cudaGetDevice(&device));
    if (device != ep->device)
        return fail(EPP_ERR_INVALID_ARG, "the current device is not the one the world was committed on");
    long long *scratch = nullptr;
    rc = static_first_row(device, &scratch);
    if (rc) return rc;
    cudaError_t e = launch_poly_trajectories(launch_cfg(*ep), ep->view, ep->bp, d_coeffs, d_durations, d_dt, n_traj,
                                             n_segments, n_coeffs, min_distance, max_samples, d_traj_valid,
                                             d_sample_mask, d_n_samples, d_first_invalid /* nullable */, scratch,
                                             (cudaStream_t)stream);
    return e == cudaSuccess ? EPP_OK : cuda_fail(e, "kernel launch");
}
