// kernels.h — launcher interface between the C-ABI layer (capi.cu) and the kernels (kernels.cu).
#pragma once

#include <cuda_runtime.h>

#include <atomic>
#include <cstdint>

#include "device_checks.cuh"

namespace epp {

struct LaunchCfg {
    int sm_count = 148;
    int states_blocks_per_sm = 8;   // 0: one block per tile (non-persistent)
    int segments_blocks_per_sm = 4; // 0: one block per tile
    int segments_variant = 0;       // 0: auto (warp-pipelined from segments_warp_min_n segments up, else direct), 1: direct with records in global memory, 2: warp-pipelined, 3: direct forms only
    int segments_warp_min_n = 32768;
    int segments_warp_threads = 1024;
    int segments_warp_slots = 0;    // cap on the ring slots per block (0: as many as fit, <= 64)
    int states_variant = 0;         // 0: auto (warp-pipelined from states_warp_min_n states up), 1: direct, 3: warp-pipelined
    int states_warp_blocks_per_sm = 1;
    int states_warp_threads = 1024; // block size of the warp-pipelined form
    int states_warp_log2_slots = 0; // cap on log2(ring slots per block); 0: as many as fit (<= 64)
    int states_warp_min_n = 4096;   // auto: batches from this size up take the warp-pipelined form
    int states_debug = 0;           // profiling only: 1 skips phase 2 of the warp form, 2 also phase 1 (wrong masks)
    int segments_block = 0;         // k_segments_rs: threads per block (256, 512, 1024; 0 = by batch size)
    int traj_variant = 0;           // pre-sampled trajectories: 0 auto (warp-pipelined form), 1 direct form only
    int states_multi = 0;           // set by the multi-batch entry point: resident blocks per SM for its launches (0: not a multi-batch launch)
    int states_multi_grid = 0;      // multi-batch launches: blocks per SM in the GRID (0: as many as are resident)
    int states_late_wait = 0;       // set per launch by the multi-batch entry point: this batch is independent of the previous launch
    int states_pdl = 1;             // warp form: programmatic dependent launch (prologue overlaps the previous kernel's tail)
    int states_bits1 = 1;           // warp form: single-word candidate masks for worlds of <= 32 OBBs
    int states_rec_smem = 1;        // warp-pipelined form: fold the world records into shared memory
    int states_rec_smem_max = 0;    // ... for worlds of up to this many OBBs (0: 256)
    int segments_multi = 0;         // K2 warp form, set per launch by the multi-batch entry point: resident blocks per SM (0 / 1: one 1024-thread block)
    int segments_late_wait = 0;     // K2 warp form, set per launch: this batch does not wait for the previous launch before streaming
    int poly_threads = 0;           // K3b: threads per block (0: by batch size, 256 x 4 or 224 x 5 per SM; 256 / 224 / 192 / 512 force a shape)
    int poly_traj_per_warp = 0;     // K3b: trajectories per block (0: by batch size)
    int poly_balance = 1;           // K3b: 1 = trajectories per block chosen to balance the SMs (see launch_poly_trajectories)
    int poly_ring = 0;              // K3b: rounds of parked samples per block (2 .. 8; 0 = 4, or 6 with 64-register threads)
    int poly_blocks_per_sm = 0;     // K3b: 4 = 256-thread blocks of 64-register threads (4 resident), 5 = never those, 0 = by batch size
    int poly_wait = 0;              // K3b: consumers sleep on the round's barrier (1) or poll with nanosleep (0)
    int gather_group_chunks = 8;    // fused exchange: chunks whose mask words travel to a peer in one run (1: each chunk alone)
    int gather_bulk_min_words = 65536; // slots from this size up: the slot is copied straight into the peers' gathered buffers
    int gather_debug = 0;           // profiling only, see GatherView::debug
    int states_per_thread = 0;      // warp form: mask words (32 states) per chunk and warp step, 4 (default) or 8
};

extern std::atomic<uint64_t> g_launches;

// Mask exchange between the GPUs of one box, done by the check kernels themselves.  Every
// rank owns a "gathered" buffer of world x slot_words mask words plus a staging area the peers write into.
// A gathering check sends every mask word it produces to every peer as an 8-byte (word, sequence number)
// pair — one store over NVLink per peer, the flag travels WITH the data, so no fence and no signal follow —
// and each of its warps, when it has run out of states, takes a share of the peers' slots: it polls the
// staging pairs until they carry this call's sequence number and writes the words into the local gathered
// buffer.  When the kernel has ended, the local gathered buffer holds the words of all ranks: an all-gather
// with no collective launch, overlapped with the checks chunk by chunk.
// Staging: 2 parities (sequence number & 1) x world source ranks x (1 + ceil(slot_words / 3)) QUADS of 16 bytes
// (three mask words + a 16-bit call flag in each 8-byte half); quad 0 of a row is the header (number of words the
// source sends).
constexpr int kMaxRanks = 8;
struct GatherView {
    uint4 *peer_ll[kMaxRanks]; // rank p's staging area in this process's address space
    const uint4 *my_ll;        // this rank's staging area
    uint64_t row_quads;        // quads per staging row: header + acknowledgement + ceil(slot_words / 3)
    uint32_t *out;             // this rank's gathered buffer of this call (world x slot_words words)
    uint32_t *peer_out[kMaxRanks]; // the peers' gathered buffers of this call (bulk form: written directly)
    uint32_t *ctl;             // local: [0] blocks done, [1] calls completed (sequence number), [2] wait time-outs, [3] bulk form: blocks that have stored
    uint64_t slot_words;
    uint64_t my_words;         // words this rank sends in this call
    uint64_t timeout_ns;       // bound on every wait (a peer that never calls must not hang the GPU)
    int world, rank;
    int debug; // profiling only (tuning knob gather_debug): 1 no wait/unpack, 2 no peer stores (wrong results)
    int deferred;       // deferred form: this call unpacks the PREVIOUS call's words (k_gather_deferred)
    uint32_t *prev_out; // ... into that call's gathered buffer (null: nothing to unpack)
};

// Request / answer lines of the single-unit mailbox (k_mailbox), in mapped pinned host memory.  The request is ONE
// 64-byte line whose two 32-byte halves each carry the sequence number (the host writes the payload first, then both
// copies; the device accepts the line when they agree, whatever the order its two halves were read in).
struct alignas(64) MailboxSlot {
    // request: [seq | flags << 32][a.x][a.y][a.z] | [b.x][b.y][b.z][seq]
    volatile uint64_t req[8];
    // answer: (seq, valid) in one 8-byte store
    volatile uint32_t ack[2];
    uint32_t pad0[14];
    // life cycle of the resident warp
    volatile uint32_t exited[2]; // (last sequence number seen, 1) once the warp has left
    volatile uint32_t state;     // kMailboxRunning / Leaving / Exited
    uint32_t pad1[13];
};
static_assert(sizeof(MailboxSlot) == 192, "MailboxSlot layout");
constexpr uint32_t kMailboxSegment = 1u, kMailboxCanPassGate = 2u, kMailboxQuit = 4u;
constexpr uint32_t kMailboxIdle = 0u, kMailboxRunning = 1u, kMailboxLeaving = 2u, kMailboxExited = 3u;
// One resident warp serving `slot` for this world until idle_ns pass without a request or a QUIT request arrives.
cudaError_t launch_mailbox(const WorldView &wv, int bp, MailboxSlot *slot, uint32_t last_seq, uint64_t idle_ns,
                           cudaStream_t stream);

cudaError_t launch_states(const LaunchCfg &cfg, const WorldView &wv, int bp, const double *d_xyz, size_t n, int mode,
                          int can_pass_gate, double min_distance, uint32_t *d_mask, cudaStream_t stream,
                          const GatherView *gv = nullptr);
cudaError_t launch_segments(const LaunchCfg &cfg, const WorldView &wv, int bp, const double *d_a, const double *d_b,
                            size_t n, int can_pass_gate, uint32_t *d_mask, cudaStream_t stream,
                            const GatherView *gv = nullptr);
// Send words [0, gv.my_words) of this rank's slot of the local gathered buffer to every peer, unpack theirs.
cudaError_t launch_gather_push(const LaunchCfg &cfg, const GatherView &gv, cudaStream_t stream);
// Deferred form: flush = 0 sends this call's words and unpacks the previous call's, flush = 1 only unpacks the last call's.
cudaError_t launch_gather_deferred(const LaunchCfg &cfg, const GatherView &gv, int flush, cudaStream_t stream);
cudaError_t launch_trajectories(const LaunchCfg &cfg, const WorldView &wv, int bp, const double *d_mats, size_t n_traj,
                                size_t rows, size_t ld, size_t traj_stride, double min_distance, uint32_t *d_traj_valid,
                                uint32_t *d_sample_mask, int64_t *d_first_invalid, long long *d_first_scratch,
                                cudaStream_t stream);
cudaError_t launch_poly_trajectories(const LaunchCfg &cfg, const WorldView &wv, int bp, const double *d_coeffs,
                                     const double *d_durations, const double *d_dt, size_t n_traj, int n_segments,
                                     int n_coeffs, double min_distance, size_t max_samples, uint32_t *d_traj_valid,
                                     uint32_t *d_sample_mask, uint32_t *d_n_samples, int64_t *d_first_invalid,
                                     long long *d_first_scratch, cudaStream_t stream);
// Batched linear min-snap solve (minsnap.cu): waypoints [n][K+1][3], durations [n][K], optional start velocity /
// acceleration [n][6] -> coefficients [n][K][3][10].
cudaError_t launch_minsnap(const double *d_waypoints, const double *d_durations, const double *d_start_vel_acc,
                           size_t n_traj, int n_segments, double *d_coeffs, cudaStream_t stream);
// The sampling half of poly_traj::generateTrajectory: column-major rows x 10 matrices [px vx ax py vy ay pz vz az t].
cudaError_t launch_sample_poly(const double *d_coeffs, const double *d_durations, const double *d_dt,
                               const double *d_time_offsets, size_t n_traj, int n_segments, int n_coeffs, size_t ld,
                               double *d_out, uint32_t *d_n_rows, cudaStream_t stream);
cudaError_t launch_gen_states(uint64_t seed, uint64_t first, size_t n, const double lo[3], const double hi[3],
                              double *d_xyz, cudaStream_t stream);
cudaError_t launch_gen_segments(uint64_t seed, uint64_t first, size_t n, const double lo[3], const double hi[3],
                                double max_len, double *d_a, double *d_b, cudaStream_t stream);
// d_out[0] = broad-phase candidates visited over the batch, d_out[1] = those passing the exact stored-AABB test
// (d_b == nullptr: points, else segments).
cudaError_t launch_count_candidates(const LaunchCfg &cfg, const WorldView &wv, int bp, const double *d_a, const double *d_b,
                                    size_t n, uint64_t *d_out, cudaStream_t stream);
cudaError_t launch_mask_summary(const LaunchCfg &cfg, const uint32_t *d_mask, size_t n_words, uint64_t first_word,
                                uint64_t *d_out, cudaStream_t stream);

} // namespace epp
