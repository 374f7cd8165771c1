/*
 * epp_oracle.h — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * Plain-C restatement of the reference planner's validity-check path
 * (Autonomous-Drone-Racing-Lab/Efficient-Path-Planner).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library; the product (libepp_b200.so) never does.
 *
 * PARITY STATUS: the reference ships no test, golden vector or fixture that
 * pins any validity result and its sources need Eigen,
 * Boost.Geometry, OMPL and yaml-cpp, none of which exist on this box.  The
 * restatement is pinned two ways instead (see DESIGN.md "Oracle"):
 *   1. hand-derived known-answer vectors (tests/golden/kat_small_portal.json)
 *   2. oracle/_ref: the reference's OWN src/OBB.cpp, src/Object.cpp,
 *      src/World.cpp, src/StateValidator.cpp, src/MotionValidator.cpp compiled
 *      unmodified, from where they lie, against stand-in headers for the absent
 *      third-party libraries (oracle/shim/).  Third-party arithmetic (Eigen
 *      product order, Boost within/intersects) is restated in the shim, so
 *      parity is "pinned against the reference's own sources, third-party
 *      arithmetic restated".
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference).  Build: gcc -O2 -ffp-contract=off (no FMA contraction, no
 * -ffast-math) — see oracle/Makefile.
 */
#ifndef EPP_ORACLE_H
#define EPP_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* One box of a component description: config.json component_geometry.<comp>.<obb>
 * (src/ConfigParserYAML.cpp:54-73; halfSize = size / 2 at :63). */
typedef struct {
    double position[3];
    double size[3];
    int filling; /* type == "filling" (1) or "collision" (0), include/OBB.h:54-57 */
} orc_obb_desc;

/* Flattened OBB record as stored in the World (include/OBB.h:47-57 plus the
 * R-tree entry of src/World.cpp:57-67). */
typedef struct {
    double center[3];
    double half[3];
    double rot[9]; /* row-major R(r,c) = rot[3*r+c] */
    double aabb_min[3];
    double aabb_max[3];
    double inflate; /* inflate radius of the owning kind (gate/obstacle) */
    int filling;
    int is_gate;
    int object_id;
    int obb_index;
} orc_obb;

typedef struct orc_world orc_world;

enum {
    ORC_OK = 0,
    ORC_ERR_ROT_X = 1,    /* "Rotation around x axis is not supported"  src/Object.cpp:38-42 */
    ORC_ERR_ROT_Y = 2,    /* "Rotation around y axis is not supported"  src/Object.cpp:43-47 */
    ORC_ERR_CENTER_Z = 3, /* "Center z position must be zero"           src/Object.cpp:16-20 */
    ORC_ERR_NO_GEOMETRY = 4, /* map::at / yaml lookup failure            src/ConfigParserYAML.cpp:21-37 */
    ORC_ERR_NO_SUCH_OBB = 5  /* obbs.at() out_of_range                   src/World.cpp:74 */
};

/* include/World.h:20-24 */
orc_world *orc_world_create(double inflate_gate, double inflate_obstacle);
void orc_world_destroy(orc_world *w);
/* gate_id_to_name_mapping + component_geometry (src/ConfigParserYAML.cpp:21-37) */
int orc_world_set_gate_geometry(orc_world *w, int type_id, const orc_obb_desc *d, int n);
int orc_world_set_obstacle_geometry(orc_world *w, const orc_obb_desc *d, int n);
/* include/World.h:29-33 */
void orc_world_reset(orc_world *w);
/* src/World.cpp:13-45 */
int orc_world_add_gate(orc_world *w, int gate_id, const double pose7[7]);
int orc_world_update_gate(orc_world *w, int gate_id, const double pose7[7]);
/* src/World.cpp:47-55 */
int orc_world_add_obstacle(orc_world *w, int obstacle_id, const double pose6[6]);
/* src/PathPlanner.cpp:60-78 */
int orc_world_parse_gates_and_obstacles(orc_world *w, const double *gates7, int n_gates,
                                        const double *obstacles6, int n_obstacles);
int orc_world_size(const orc_world *w);
int orc_world_get(const orc_world *w, int i, orc_obb *out);

/* src/OBB.cpp:63-91 / :10-61 / :93-124 on a record */
int orc_obb_point_hit(const orc_obb *o, const double p[3], double inflate);
int orc_obb_ray_hit(const orc_obb *o, const double a[3], const double b[3], double inflate);

/* src/World.cpp:80-104, :106-128, :130-163 — return 1 = valid */
int orc_check_point(const orc_world *w, const double p[3], int can_pass_gate);
int orc_check_point_mindist(const orc_world *w, const double p[3], double min_distance);
int orc_check_ray(const orc_world *w, const double a[3], const double b[3], int can_pass_gate);

/* Batched loops over the predicates.  mask bit (i&31) of word (i>>5) = 1 iff
 * valid; tail bits zero.  threads <= 1 runs on the caller's thread. */
void orc_check_states(const orc_world *w, const double *xyz, size_t n, int can_pass_gate,
                      uint32_t *mask, int threads);
void orc_check_states_mindist(const orc_world *w, const double *xyz, size_t n, double min_distance,
                              uint32_t *mask, int threads);
void orc_check_segments(const orc_world *w, const double *a, const double *b, size_t n,
                        int can_pass_gate, uint32_t *mask, int threads);

/* src/PathPlanner.cpp:267-280 on a column-major rows x cols matrix (position
 * columns 0,3,6).  Returns 1 = valid.  sample_mask (optional) gets one bit per
 * row evaluated WITHOUT the early exit; first_invalid (optional) gets the first
 * invalid row or -1. */
int orc_check_trajectory(const orc_world *w, const double *col_major, size_t rows, size_t ld,
                         double min_distance, uint32_t *sample_mask, long *first_invalid);

/* external/poly_traj/include/mav_trajectory_generation/polynomial.h:136-150 (derivative 0) */
double orc_poly_eval(const double *coeffs, int n_coeffs, double t);
/* external/poly_traj/src/trajectory.cpp:81-141 with derivative 0 and
 * external/poly_traj/src/segment.cpp:51-58.  coeffs[(k*3+d)*n_coeffs + j],
 * durations[k].  Emits up to max_out samples; returns the number the reference
 * would emit (may exceed max_out). */
size_t orc_traj_evaluate_range(int n_segments, int n_coeffs, const double *coeffs,
                               const double *durations, double t_start, double t_end, double dt,
                               double *out_xyz, double *out_times, size_t max_out);
/* external/poly_traj/include/mav_trajectory_generation/trajectory.h:64-71 */
double orc_traj_max_time(int n_segments, const double *durations);
/* Fused sweep = evaluateRange(0, maxTime, dt) then checkTrajectoryValidity rows.
 * Returns 1 = valid; n_samples_out = sample count. */
int orc_check_poly_traj(const orc_world *w, int n_segments, int n_coeffs, const double *coeffs,
                        const double *durations, double dt, double min_distance,
                        uint32_t *sample_mask, size_t mask_words, size_t *n_samples_out,
                        long *first_invalid);

/* Counter-based synthetic input generators shared (bit-identically) with the
 * device generators in the product library (DESIGN.md "Synthetic inputs"). */
uint64_t orc_mix64(uint64_t seed, uint64_t counter);
void orc_gen_states(uint64_t seed, uint64_t first, size_t n, const double lo[3], const double hi[3],
                    double *xyz);
void orc_gen_segments(uint64_t seed, uint64_t first, size_t n, const double lo[3],
                      const double hi[3], double max_len, double *a, double *b);

#ifdef __cplusplus
}
#endif
#endif
