/*
 * epp_c_api.h — C ABI of libepp_b200.so, the B200-native (sm_100a) replacement for the
 * validity-check hot path of Autonomous-Drone-Racing-Lab/Efficient-Path-Planner.
 *
 * Every entry point names the reference interface it replaces (paths relative to the
 * reference repository root).  Plain pointers and sizes only; no C++/torch types cross
 * this boundary.  All functions return an epp_status (0 = ok); the message of the last
 * failure on the calling thread is available from epp_last_error().  There is no CPU
 * fallback: every check runs on the GPU and fails with EPP_ERR_NO_DEVICE / EPP_ERR_CUDA
 * when it cannot.
 *
 * Mask convention (all batched checks): bit (i & 31) of word (i >> 5) is 1 iff unit i is
 * VALID (collision free); tail bits of the last word are 0.  A batch of n units needs
 * (n + 31) / 32 uint32 words.
 *
 * Inputs: coordinates must be finite.  The kernels use the z-rotation structure of every world this API can build
 * (src/Object.cpp:38-47 rejects roll and pitch): the exact 0 * d terms of the reference's 3-term sums are dropped, which
 * is bit-identical for finite inputs and undefined for NaN / infinite ones (the reference itself returns garbage there).
 *
 * Device-pointer entry points (*_dev) only enqueue work on `stream`; they may be captured into CUDA graphs.  A captured
 * graph reads the world epoch that was current at capture time: re-capture after epp_world_commit (the retired epoch's
 * memory is released once no call in flight holds it).
 *
 * Threading: check calls are re-entrant and may be issued concurrently from several host
 * threads on one world (the reference runs two planner threads on one World,
 * src/OnlineTrajGenerator.cpp:324-340).  World mutation (reset/add/update/commit) must be
 * externally serialised against other mutations; checks in flight keep using the epoch
 * they started with.
 */
#ifndef EPP_C_API_H
#define EPP_C_API_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct epp_world epp_world;

typedef enum {
    EPP_OK = 0,
    EPP_ERR_ROT_X = 1,       /* "Rotation around x axis is not supported"  src/Object.cpp:38-42 */
    EPP_ERR_ROT_Y = 2,       /* "Rotation around y axis is not supported"  src/Object.cpp:43-47 */
    EPP_ERR_CENTER_Z = 3,    /* "Center z position must be zero"           src/Object.cpp:16-20 */
    EPP_ERR_NO_GEOMETRY = 4, /* unknown gate type / component              src/ConfigParserYAML.cpp:21-37 */
    EPP_ERR_NO_SUCH_OBB = 5, /* std::out_of_range from obbs.at()           src/World.cpp:74,87 */
    EPP_ERR_INVALID_ARG = 6,
    EPP_ERR_NOT_COMMITTED = 7, /* check issued before the first epp_world_commit */
    EPP_ERR_CUDA = 8,
    EPP_ERR_NO_DEVICE = 9
} epp_status;

/* One box of a component description: config component_geometry.<comp>.<obb>
 * (include/Types.h:22-28 OBBDescription; halfSize = size / 2, src/ConfigParserYAML.cpp:63). */
typedef struct {
    double position[3];
    double size[3];
    int filling; /* 1: type "filling", 0: type "collision" (include/OBB.h:54-57) */
} epp_obb_desc;

/* Host copy of one world record (include/OBB.h:47-57 + its R-tree value, src/World.cpp:57-67). */
typedef struct {
    double center[3];
    double half[3];
    double rot[9]; /* row-major */
    double aabb_min[3];
    double aabb_max[3];
    double inflate;
    int filling;
    int is_gate;
    int object_id;
    int obb_index;
} epp_obb;

/* Broad-phase selection (DESIGN.md "Broad phase").  Results are identical for all modes. */
typedef enum {
    EPP_BP_AUTO = 0,
    EPP_BP_BRUTE = 1, /* every OBB's stored AABB tested per unit */
    EPP_BP_BITS = 2,  /* per-axis bin bitmasks, M <= 1024 */
    EPP_BP_GRID = 3   /* uniform 3-D grid with per-cell candidate lists, any M */
} epp_broadphase;

const char *epp_last_error(void);
const char *epp_version(void);
int epp_device_count(int *count);

/* ---- World lifecycle ---------------------------------------------------------------- */

/* World::World(configParser): inflate radii per kind (include/World.h:20-24). */
int epp_world_create(double inflate_gate, double inflate_obstacle, epp_world **out);
int epp_world_destroy(epp_world *w);
/* ConfigParser::getGateGeometryByTypeId / getObstacleGeometry (src/ConfigParserYAML.cpp:21-37). */
int epp_world_set_gate_geometry(epp_world *w, int type_id, const epp_obb_desc *descs, int n);
int epp_world_set_obstacle_geometry(epp_world *w, const epp_obb_desc *descs, int n);
/* World::resetWorld (include/World.h:29-33). */
int epp_world_reset(epp_world *w);
/* World::addGate(gateId, [x,y,z,roll,pitch,yaw,type]) (src/World.cpp:13-40). */
int epp_world_add_gate(epp_world *w, int gate_id, const double pose7[7]);
/* World::updateGatePosition (src/World.cpp:42-45, removeObject :69-78). */
int epp_world_update_gate(epp_world *w, int gate_id, const double pose7[7]);
/* World::addObstacle(obstacleId, [x,y,z,roll,pitch,yaw]) (src/World.cpp:47-55). */
int epp_world_add_obstacle(epp_world *w, int obstacle_id, const double pose6[6]);
/* PathPlanner::parseGatesAndObstacles (src/PathPlanner.cpp:60-78): reset, gates (z forced
 * to 0) then obstacles, row-major n x 7 / n x 6, followed by a commit. */
int epp_world_parse_gates_and_obstacles(epp_world *w, const double *gates7, int n_gates,
                                        const double *obstacles6, int n_obstacles);
int epp_world_size(const epp_world *w, int *n_obbs);
int epp_world_get_obb(const epp_world *w, int index, epp_obb *out);
int epp_world_set_broadphase(epp_world *w, epp_broadphase mode);
/* Publishes the current host world as a new immutable device epoch on the calling
 * thread's current CUDA device (replaces the reference's in-place R-tree mutation,
 * src/World.cpp:57-78).  Checks see the world as of the LAST commit: reset/add/update stay invisible to
 * them (also to checks other threads start meanwhile) until the next commit. */
int epp_world_commit(epp_world *w);
int epp_world_epoch(const epp_world *w, uint64_t *epoch);
/* Broad phase actually in use for the committed epoch (an epp_broadphase, never AUTO). */
int epp_world_active_broadphase(const epp_world *w, int *mode);

/* ---- Checks on HOST buffers (H2D copy, kernel, D2H copy; synchronous) ----------------- */

/* World::checkPointValidity(point, bool canPassGate) over n points (src/World.cpp:80-104);
 * xyz is n x 3 row-major. */
int epp_check_states(epp_world *w, const double *xyz, size_t n, int can_pass_gate, uint32_t *mask);
/* World::checkPointValidity(point, double minDistance) over n points (src/World.cpp:106-128). */
int epp_check_states_mindist(epp_world *w, const double *xyz, size_t n, double min_distance,
                             uint32_t *mask);
/* World::checkRayValid(start, end, canPassGate) over n segments (src/World.cpp:130-163). */
int epp_check_segments(epp_world *w, const double *a, const double *b, size_t n, int can_pass_gate,
                       uint32_t *mask);
/* StateValidator::isValid for one state (src/StateValidator.cpp:7-13). */
int epp_is_valid(epp_world *w, const double p[3], int can_pass_gate, int *valid);
/* MotionValidator::checkMotion(s1, s2) for one edge (src/MotionValidator.cpp:8-17). */
int epp_check_motion(epp_world *w, const double a[3], const double b[3], int can_pass_gate,
                     int *valid);
/* PathPlanner::checkTrajectoryValidity(trajectory, minDistance) (src/PathPlanner.cpp:267-280)
 * for n_traj column-major matrices of `rows` rows and leading dimension `ld`, consecutive
 * matrices `traj_stride` doubles apart; position = columns 0, 3, 6 (so cols >= 7).
 * traj_valid: one bit per trajectory.  sample_mask (nullable): per trajectory
 * (rows + 31) / 32 words, one bit per row (no early exit).  first_invalid (nullable): per
 * trajectory the first invalid row or -1. */
int epp_check_trajectories(epp_world *w, const double *col_major, size_t n_traj, size_t rows,
                           size_t ld, size_t traj_stride, int cols, double min_distance,
                           uint32_t *traj_valid, uint32_t *sample_mask, int64_t *first_invalid);
/* Fused sweep: Trajectory::evaluateRange(0, maxTime, dt) (external/poly_traj/src/trajectory.cpp:
 * 81-141) + Segment::evaluate (external/poly_traj/src/segment.cpp:51-58) + Polynomial::evaluate
 * (external/poly_traj/include/mav_trajectory_generation/polynomial.h:136-150) feeding
 * checkTrajectoryValidity, for n_traj trajectories of n_segments segments each.
 * coeffs[((t * n_segments + k) * 3 + d) * n_coeffs + j], durations[t * n_segments + k],
 * dt[t] (per-trajectory sampling interval).  max_samples bounds the samples per trajectory
 * (multiple of 32); n_samples (nullable) returns the count the sampler produced. */
int epp_check_poly_trajectories(epp_world *w, const double *coeffs, const double *durations,
                                const double *dt, size_t n_traj, int n_segments, int n_coeffs,
                                double min_distance, size_t max_samples, uint32_t *traj_valid,
                                uint32_t *sample_mask, uint32_t *n_samples, int64_t *first_invalid);

/* ---- Checks on DEVICE buffers (asynchronous on `stream`, a cudaStream_t; capture-safe) - */

int epp_check_states_dev(epp_world *w, const double *d_xyz, size_t n, int can_pass_gate,
                         uint32_t *d_mask, void *stream);
/* Several INDEPENDENT state batches in one call (the added batched entry point at the size the planner produces:
 * many sample batches per solve): batch i = n[i] states at d_xyz[i] -> mask words at d_mask[i].  Launched back to
 * back on `stream`; from the second batch on a kernel starts streaming its states while the previous one is in its
 * tail.  Contract: all inputs are complete before the call and no batch's mask overlaps another batch's buffers
 * (overlapping masks are refused). */
int epp_check_states_multi_dev(epp_world *w, const double *const *d_xyz, const size_t *n, uint32_t *const *d_mask,
                               int batches, int can_pass_gate, void *stream);
int epp_check_states_mindist_dev(epp_world *w, const double *d_xyz, size_t n, double min_distance,
                                 uint32_t *d_mask, void *stream);
int epp_check_segments_dev(epp_world *w, const double *d_a, const double *d_b, size_t n,
                           int can_pass_gate, uint32_t *d_mask, void *stream);
/* `batches` independent segment batches in one call (the edge counterpart of epp_check_states_multi_dev: a planner's
 * batched MotionValidator::checkMotion, src/MotionValidator.cpp:9-17 -> World::checkRayValid, src/World.cpp:130-163).
 * Same results as `batches` calls of epp_check_segments_dev; because the library knows the batches are independent
 * and their inputs complete, a batch streams its end points while the previous one is still in its tail.  The masks
 * must not overlap each other (EPP_ERR_INVALID_ARG); n[i] may be 0. */
int epp_check_segments_multi_dev(epp_world *w, const double *const *d_a, const double *const *d_b, const size_t *n,
                                 uint32_t *const *d_mask, int batches, int can_pass_gate, void *stream);
/* d_first_invalid (nullable) is int64 per trajectory; d_traj_valid words must be zeroed by
 * the callee (they are). */
int epp_check_trajectories_dev(epp_world *w, const double *d_col_major, size_t n_traj, size_t rows,
                               size_t ld, size_t traj_stride, int cols, double min_distance,
                               uint32_t *d_traj_valid, uint32_t *d_sample_mask,
                               int64_t *d_first_invalid, void *stream);
int epp_check_poly_trajectories_dev(epp_world *w, const double *d_coeffs, const double *d_durations,
                                    const double *d_dt, size_t n_traj, int n_segments, int n_coeffs,
                                    double min_distance, size_t max_samples, uint32_t *d_traj_valid,
                                    uint32_t *d_sample_mask, uint32_t *d_n_samples,
                                    int64_t *d_first_invalid, void *stream);
/* Allocation-free variant of epp_check_poly_trajectories_dev.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same arguments, same results bit for bit, except that
 * d_first_invalid may be NULL: the per-trajectory "first invalid row" words the kernels work on (which the entry
 * point above takes from the caller's d_first_invalid) live in a static device array of
 * EPP_STATIC_MAX_TRAJECTORIES int64 words inside the library (512 KiB on every GPU that loads it, until exit).
 *  - n_traj > EPP_STATIC_MAX_TRAJECTORIES returns EPP_ERR_INVALID_ARG (epp_last_error names the limit); nothing is
 *    launched.  Split the batch or use the entry point above.
 *  - The call makes no cudaMalloc / cudaFree and calls no allocating library routine; the world must already be
 *    committed (epp_world_commit allocates, this does not) on the CURRENT device, else EPP_ERR_NOT_COMMITTED /
 *    EPP_ERR_INVALID_ARG.
 *  - Every call shares the array, which keeps the last call's contents; the words a call uses are reset on its
 *    stream before its kernels read them.  Calls that OVERLAP on the GPU (different streams or threads) race on
 *    it: serialise them. */
#define EPP_STATIC_MAX_TRAJECTORIES 65536
int epp_check_poly_trajectories_static_dev(epp_world *w, const double *d_coeffs, const double *d_durations,
                                           const double *d_dt, size_t n_traj, int n_segments, int n_coeffs,
                                           double min_distance, size_t max_samples, uint32_t *d_traj_valid,
                                           uint32_t *d_sample_mask, uint32_t *d_n_samples,
                                           int64_t *d_first_invalid /* nullable */, void *stream);

/* ---- Min-snap trajectory generation, batched --------------------------
 *
 * Replaces, for many waypoint sets at once, the solve inside poly_traj::generateTrajectory
 * (external/poly_traj/src/trajectory_generator.cpp:12-58): vertices with the start fully
 * constrained (position, velocity, acceleration given; jerk = snap = 0), interior vertices
 * position-only, the end at rest; PolynomialOptimization<10>::setupFromVertices + solveLinear
 * (impl/polynomial_optimization_linear_impl.h:56-380), derivative_to_optimize = SNAP.
 * Output layout = input layout of epp_check_poly_trajectories: coeffs[n][K][3][10] (increasing
 * powers).  Floating-point result: agrees with the reference formulation to ~1e-9 m along the
 * trajectory (tests/test_minsnap.py), not bit for bit. */
/* estimateSegmentTimes = estimateSegmentTimesNfabian (vertex.cpp:245-289), host side (libm exp):
 * waypoints[n_vertices][3] -> durations[n_vertices - 1]. */
int epp_segment_times(const double *waypoints, int n_vertices, double v_max, double a_max, double *durations);
/* start_vel_acc: n_traj x 6 (velocity xyz, acceleration xyz) or NULL for a start at rest. */
int epp_minsnap(const double *waypoints, const double *durations, const double *start_vel_acc, size_t n_traj,
                int n_segments, double *coeffs);
int epp_minsnap_dev(const double *d_waypoints, const double *d_durations, const double *d_start_vel_acc,
                    size_t n_traj, int n_segments, double *d_coeffs, void *stream);

/* The sampling half of poly_traj::generateTrajectory (trajectory_generator.cpp:60-97): for every trajectory
 * the rows x 10 matrix [px vx ax py vy ay pz vz az t + offset] of evaluateRange(0, maxTime, dt) in
 * position, velocity and acceleration — bit-equal to the reference's — column-major with leading
 * dimension ld (rows beyond ld are counted in n_rows but not written), matrices ld * 10 doubles apart. */
int epp_sample_trajectories_dev(const double *d_coeffs, const double *d_durations, const double *d_dt,
                                const double *d_time_offsets /* nullable */, size_t n_traj, int n_segments,
                                int n_coeffs, size_t ld, double *d_out, uint32_t *d_n_rows, void *stream);
int epp_sample_trajectories(const double *coeffs, const double *durations, const double *dt,
                            const double *time_offsets /* nullable */, size_t n_traj, int n_segments, int n_coeffs,
                            size_t ld, double *out, uint32_t *n_rows);

/* ---- Multi-GPU sharding helpers (host logic) --------------------------- */

/* Contiguous range of units for `rank` of `world_size`, boundaries aligned to `align`
 * units (a multiple of 32 so every rank owns whole mask words). */
int epp_shard_range(size_t n, int rank, int world_size, size_t align, size_t *begin, size_t *end);
/* popcount and 64-bit XOR-fold checksum of a device mask (words consecutive pairs). */
int epp_mask_summary_dev(const uint32_t *d_mask, size_t n_words, uint64_t first_word, uint64_t *d_popcount_xor,
                         void *stream);

/* ---- Mask exchange fused into the check kernels --------------------------
 *
 * One epp_gather per rank (= per GPU; one process per GPU, or several handles in one process).
 * It owns `n_buffers` gathered buffers of world_size x slot_words mask words in device memory.
 * A gathering check sends every mask word it produces to every peer as an 8-byte
 * (word, call number) pair — one NVLink store per peer, straight from the check kernel, the
 * flag travelling with the data so that no fence, signal or collective launch follows — and
 * the kernel's warps, once out of units, unpack the pairs the peers sent into slot p of the
 * local buffer.  When the call has completed in stream order, epp_gather_buffer(g, buffer)
 * holds the all-gathered mask (words beyond a rank's units are 0).  Gathering calls are
 * collective: every rank issues the same sequence of them, each on one stream; the ranks'
 * kernels wait for each other on the device, so they must be able to run concurrently (one
 * GPU each).  A peer that does not show up within the time-out (default 2 s) is counted in
 * epp_gather_status and read as 0 — the GPU never hangs. */
typedef struct epp_gather epp_gather;
#define EPP_GATHER_HANDLE_BYTES 64
#define EPP_GATHER_MAX_RANKS 8

int epp_gather_create(int rank, int world_size, size_t slot_words, int n_buffers, epp_gather **out);
int epp_gather_destroy(epp_gather *g);
/* Inter-process: export this rank's allocation (a CUDA IPC handle), exchange the handles by
 * any means (torch.distributed, MPI, a file), connect with all of them in rank order. */
int epp_gather_export(epp_gather *g, unsigned char handle[EPP_GATHER_HANDLE_BYTES]);
int epp_gather_connect(epp_gather *g, const unsigned char *handles /* world_size x 64 B */);
/* In-process (one host process driving several GPUs, or several ranks on one GPU in tests):
 * connects all handles with each other; enables peer access between their devices. */
int epp_gather_connect_local(epp_gather *const *all, int world_size);
/* Device pointer of gathered buffer `buffer` (world_size x slot_words words, rank-major). */
int epp_gather_buffer(epp_gather *g, int buffer, uint32_t **d_words);
int epp_gather_set_timeout_ms(epp_gather *g, unsigned ms);
/* Deferred form, for steps so short that waiting for the peers dominates (1 M states per rank: a 9 us check inside a
 * 15 us step at 8 ranks): with it on, a gathering call sends its words and unpacks the peers' words of the PREVIOUS
 * call, which arrived a step ago, so nothing waits.  Buffer b of call s is complete (all ranks' words) once call
 * s + 1 or epp_gather_flush has completed in stream order; until then it holds this rank's own slot only.  Consecutive
 * calls must use different buffers (n_buffers >= 2).  Slots large enough for the bulk form ignore the setting. */
int epp_gather_set_deferred(epp_gather *g, int on);
int epp_gather_flush(epp_gather *g, void *stream);
/* Calls completed so far on this rank and how many of their waits timed out (synchronises the device). */
int epp_gather_status(epp_gather *g, uint64_t *calls, uint64_t *timeouts);

/* epp_check_states_dev / epp_check_segments_dev for this rank's n units + the exchange. */
int epp_check_states_gather_dev(epp_world *w, const double *d_xyz, size_t n, int can_pass_gate,
                                epp_gather *g, int buffer, void *stream);
int epp_check_states_mindist_gather_dev(epp_world *w, const double *d_xyz, size_t n, double min_distance,
                                        epp_gather *g, int buffer, void *stream);
int epp_check_segments_gather_dev(epp_world *w, const double *d_a, const double *d_b, size_t n,
                                  int can_pass_gate, epp_gather *g, int buffer, void *stream);

/* ---- Synthetic inputs generated on the device (counter based, bit-identical to the
 *      oracle's host generators; DESIGN.md "Synthetic inputs") -------------------------- */

int epp_gen_states_dev(uint64_t seed, uint64_t first, size_t n, const double lo[3],
                       const double hi[3], double *d_xyz, void *stream);
int epp_gen_segments_dev(uint64_t seed, uint64_t first, size_t n, const double lo[3],
                         const double hi[3], double max_len, double *d_a, double *d_b, void *stream);

/* ---- Introspection used by bench.py --------------------------------------------------- */

/* Number of kernels this library has launched on the calling process so far. */
uint64_t epp_kernel_launch_count(void);
/* Mean wall time (microseconds) of n back-to-back epp_is_valid calls, timed inside the library (no binding overhead):
 * what one StateValidator::isValid (src/StateValidator.cpp:7-13) costs an OMPL planner thread. */
int epp_measure_single_latency(epp_world *w, int n_calls, double *us_per_call);
/* Broad-phase statistics of a device batch against the committed world (d_b NULL: points, else segments):
 * d_visited_hits[0] = candidates the active broad phase visits, [1] = those whose stored AABB passes the
 * reference's exact R-tree predicate (src/World.cpp:83 strict containment / :143 closed intersection).  Feeds the
 * FP64 ops-per-check figure of the roofline (6 * M_visited + 21 * k). */
int epp_count_candidates_dev(epp_world *w, const double *d_a, const double *d_b /* nullable */, size_t n,
                             uint64_t *d_visited_hits, void *stream);
/* Non-fused FP64 issue rate of the current device (thread-level DADD + DMUL instructions per second; 8
 * independent chains per thread, full occupancy, no DFMA in the SASS): the compute roof of this path, whose
 * reference arithmetic (src/OBB.cpp:10-91, polynomial.h:136-150) is individually rounded.  Best of `reps`
 * launches of `iters` loop iterations, timed with CUDA events on `stream`. */
int epp_measure_fp64_nofma(int iters, int reps, void *stream, double *instr_per_s, double *ms_per_launch);
/* Tuning knobs (DESIGN.md "Tuning"); unknown keys return EPP_ERR_INVALID_ARG. */
int epp_set_tuning(const char *key, int value);
int epp_get_tuning(const char *key, int *value);

#ifdef __cplusplus
}
#endif
#endif
