/*
 * epp_oracle.c — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 * See epp_oracle.h for scope, parity status and build flags.
 *
 * Arithmetic rules: every + - * / is an individually
 * rounded binary64 operation (compile with -ffp-contract=off); 3-term dot
 * products are summed left to right, (x0 + x1) + x2; std::min(a,b) is
 * (b<a)?b:a and std::max(a,b) is (a<b)?b:a; unqualified abs(double) in the
 * reference resolves to the double overload, restated with fabs().
 */
#include "epp_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ types */

typedef struct {
    int n;
    orc_obb_desc *d;
} geom_list;

/* One std::map<std::string,OBB> node (include/World.h:112).  The key string
 * "{gate|obstacle}_{id}_obb_{i}" (src/World.cpp:63) is the triple below. */
typedef struct {
    int is_gate, object_id, obb_index;
    double center[3], half[3], rot[9];
    int filling;
} map_obb;

/* One R-tree value pair<box,string> (include/Types.h:13-16). */
typedef struct {
    double bmin[3], bmax[3];
    int is_gate, object_id, obb_index;
} tree_entry;

#define MAX_GATE_TYPES 64

struct orc_world {
    double inflate_gate, inflate_obstacle;
    geom_list gate_geom[MAX_GATE_TYPES];
    geom_list obstacle_geom;
    int n_map, cap_map;
    map_obb *map;
    int n_tree, cap_tree;
    tree_entry *tree;
};

/* ------------------------------------------------------- small vector math */

static inline double min_ref(double a, double b) { return (b < a) ? b : a; } /* std::min */
static inline double max_ref(double a, double b) { return (a < b) ? b : a; } /* std::max */

/* Eigen fixed-size 3x3 * 3x1 coefficient-based product: out_r = sum_c M(r,c) v_c,
 * summed left to right (SURVEY.md §8a note E). */
static void mat_vec(const double m[9], const double v[3], double out[3])
{
    for (int r = 0; r < 3; ++r)
        out[r] = (m[3 * r + 0] * v[0] + m[3 * r + 1] * v[1]) + m[3 * r + 2] * v[2];
}

/* rotation.transpose() * v : out_i = sum_k R(k,i) v_k */
static void mat_t_vec(const double m[9], const double v[3], double out[3])
{
    for (int i = 0; i < 3; ++i)
        out[i] = (m[0 + i] * v[0] + m[3 + i] * v[1]) + m[6 + i] * v[2];
}

static void mat_mat(const double a[9], const double b[9], double out[9])
{
    double t[9];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c)
            t[3 * r + c] = (a[3 * r + 0] * b[0 + c] + a[3 * r + 1] * b[3 + c]) + a[3 * r + 2] * b[6 + c];
    memcpy(out, t, sizeof t);
}

/* ------------------------------------------------------------ OBB geometry */

/* src/OBB.cpp:63-91.  `inflate_this` = shouldBeInflated() (include/OBB.h:54-57). */
static int point_hit(const double center[3], const double half[3], const double rot[9],
                     int filling, const double p[3], double inflate)
{
    double d[3] = {p[0] - center[0], p[1] - center[1], p[2] - center[2]};
    double local[3];
    mat_t_vec(rot, d, local); /* :66 */
    double h[3] = {half[0], half[1], half[2]};
    if (!filling) { /* :69-72 */
        h[0] += inflate;
        h[1] += inflate;
        h[2] += inflate;
    }
    return fabs(local[0]) <= h[0] && fabs(local[1]) <= h[1] && fabs(local[2]) <= h[2]; /* :75-77 */
}

/* src/OBB.cpp:10-61 */
static int ray_hit(const double center[3], const double half[3], const double rot[9], int filling,
                   const double a[3], const double b[3], double inflate)
{
    int hit_a = point_hit(center, half, rot, filling, a, inflate); /* :13 */
    int hit_b = point_hit(center, half, rot, filling, b, inflate); /* :14 */
    if (hit_a || hit_b) return 1;                                  /* :15-18 */

    double da[3] = {a[0] - center[0], a[1] - center[1], a[2] - center[2]};
    double db[3] = {b[0] - center[0], b[1] - center[1], b[2] - center[2]};
    double ls[3], le[3], dir[3];
    mat_t_vec(rot, da, ls); /* :21 */
    mat_t_vec(rot, db, le); /* :22 */
    for (int i = 0; i < 3; ++i) dir[i] = le[i] - ls[i]; /* :23 */

    double t_min = 0, t_max = 1; /* :25-26 */
    double bmin[3], bmax[3];
    for (int i = 0; i < 3; ++i) { /* :28-30, inflated for every type */
        double h = half[i] + inflate;
        bmin[i] = -h;
        bmax[i] = h;
    }
    for (int i = 0; i < 3; ++i) { /* :32-59 */
        if (fabs(dir[i]) < 1e-6) {
            if (ls[i] < bmin[i] || ls[i] > bmax[i]) return 0;
        } else {
            double inv = 1.0 / dir[i]; /* 1.0f promotes exactly, :44 */
            double t1 = (bmin[i] - ls[i]) * inv;
            double t2 = (bmax[i] - ls[i]) * inv;
            double t_entry = min_ref(t1, t2);
            double t_exit = max_ref(t1, t2);
            t_min = max_ref(t_min, t_entry);
            t_max = min_ref(t_max, t_exit);
            if (t_min > t_max) return 0;
        }
    }
    return 0 <= t_min && t_min <= 1 && 0 <= t_max && t_max <= 1; /* :60 */
}

/* src/OBB.cpp:93-124 */
static void obb_aabb(const double center[3], const double half[3], const double rot[9],
                     int filling, double inflate, double out_min[3], double out_max[3])
{
    static const double sx[8] = {-1, 1, 1, -1, -1, 1, 1, -1}; /* :100-102 */
    static const double sy[8] = {-1, -1, 1, 1, -1, -1, 1, 1};
    static const double sz[8] = {-1, -1, -1, -1, 1, 1, 1, 1};
    for (int c = 0; c < 8; ++c) {
        double corner[3] = {sx[c] * half[0], sy[c] * half[1], sz[c] * half[2]}; /* :104-105 */
        double g[3];
        mat_vec(rot, corner, g);
        for (int r = 0; r < 3; ++r) {
            double v = g[r] + center[r]; /* :110 */
            if (c == 0) {
                out_min[r] = v;
                out_max[r] = v;
            } else {
                if (v < out_min[r]) out_min[r] = v; /* :111-112 */
                if (v > out_max[r]) out_max[r] = v;
            }
        }
    }
    if (!filling) { /* :117-121 */
        for (int r = 0; r < 3; ++r) {
            out_min[r] -= inflate;
            out_max[r] += inflate;
        }
    }
}

int orc_obb_point_hit(const orc_obb *o, const double p[3], double inflate)
{
    return point_hit(o->center, o->half, o->rot, o->filling, p, inflate);
}

int orc_obb_ray_hit(const orc_obb *o, const double a[3], const double b[3], double inflate)
{
    return ray_hit(o->center, o->half, o->rot, o->filling, a, b, inflate);
}

/* ------------------------------------------------------------------ World */

orc_world *orc_world_create(double inflate_gate, double inflate_obstacle)
{
    orc_world *w = (orc_world *)calloc(1, sizeof *w);
    w->inflate_gate = inflate_gate;
    w->inflate_obstacle = inflate_obstacle;
    return w;
}

static void geom_free(geom_list *g)
{
    free(g->d);
    g->d = NULL;
    g->n = 0;
}

void orc_world_destroy(orc_world *w)
{
    if (!w) return;
    for (int i = 0; i < MAX_GATE_TYPES; ++i) geom_free(&w->gate_geom[i]);
    geom_free(&w->obstacle_geom);
    free(w->map);
    free(w->tree);
    free(w);
}

static void geom_set(geom_list *g, const orc_obb_desc *d, int n)
{
    geom_free(g);
    g->d = (orc_obb_desc *)malloc(sizeof(orc_obb_desc) * (size_t)(n > 0 ? n : 1));
    memcpy(g->d, d, sizeof(orc_obb_desc) * (size_t)n);
    g->n = n;
}

int orc_world_set_gate_geometry(orc_world *w, int type_id, const orc_obb_desc *d, int n)
{
    if (type_id < 0 || type_id >= MAX_GATE_TYPES) return ORC_ERR_NO_GEOMETRY;
    geom_set(&w->gate_geom[type_id], d, n);
    return ORC_OK;
}

int orc_world_set_obstacle_geometry(orc_world *w, const orc_obb_desc *d, int n)
{
    geom_set(&w->obstacle_geom, d, n);
    return ORC_OK;
}

void orc_world_reset(orc_world *w) /* include/World.h:29-33 */
{
    w->n_map = 0;
    w->n_tree = 0;
}

static map_obb *map_find(const orc_world *w, int is_gate, int id, int idx)
{
    for (int i = 0; i < w->n_map; ++i)
        if (w->map[i].is_gate == is_gate && w->map[i].object_id == id && w->map[i].obb_index == idx)
            return &w->map[i];
    return NULL;
}

/* Object::createFromDescription + Object::Object + translate + rotateZ
 * (src/Object.cpp:11-85) producing the object's OBB list. */
static int build_object(const geom_list *g, const double pos[3], const double rot[3],
                        map_obb *out)
{
    if (fabs(rot[0]) > 1e-6) return ORC_ERR_ROT_X; /* :38-42 */
    if (fabs(rot[1]) > 1e-6) return ORC_ERR_ROT_Y; /* :43-47 */
    if (pos[2] > 1e-6) return ORC_ERR_CENTER_Z;    /* :16-20 */

    double gc[3] = {0, 0, 0};
    for (int r = 0; r < 3; ++r) gc[r] += pos[r]; /* translate, :54 */

    double c = cos(rot[2]), s = sin(rot[2]); /* :69-70 */
    double rz[9] = {c, -s, 0, s, c, 0, 0, 0, 1}; /* :72-75 */

    for (int i = 0; i < g->n; ++i) {
        map_obb *o = &out[i];
        for (int r = 0; r < 3; ++r) {
            o->center[r] = g->d[i].position[r];
            o->half[r] = g->d[i].size[r] / 2; /* ConfigParserYAML.cpp:63 */
        }
        static const double ident[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}; /* OBB.h:49 */
        memcpy(o->rot, ident, sizeof ident);
        o->filling = g->d[i].filling;
        for (int r = 0; r < 3; ++r) o->center[r] += pos[r]; /* translate, :55-58 */
        double rel[3] = {o->center[0] - gc[0], o->center[1] - gc[1], o->center[2] - gc[2]}; /* :81 */
        double rc[3];
        mat_vec(rz, rel, rc);
        for (int r = 0; r < 3; ++r) o->center[r] = rc[r] + gc[r]; /* :82 */
        mat_mat(rz, o->rot, o->rot);                             /* :83 */
    }
    return ORC_OK;
}

/* World::addObject, src/World.cpp:57-67.  std::map::insert keeps an existing
 * key's OBB; the R-tree always gains the new (aabb,key) value. */
static void add_object(orc_world *w, const map_obb *objs, int n, int id, int is_gate, double inflate)
{
    for (int i = 0; i < n; ++i) {
        double mn[3], mx[3];
        obb_aabb(objs[i].center, objs[i].half, objs[i].rot, objs[i].filling, inflate, mn, mx);
        if (!map_find(w, is_gate, id, i)) {
            if (w->n_map == w->cap_map) {
                w->cap_map = w->cap_map ? 2 * w->cap_map : 64;
                w->map = (map_obb *)realloc(w->map, sizeof(map_obb) * (size_t)w->cap_map);
            }
            w->map[w->n_map] = objs[i];
            w->map[w->n_map].is_gate = is_gate;
            w->map[w->n_map].object_id = id;
            w->map[w->n_map].obb_index = i;
            w->n_map++;
        }
        if (w->n_tree == w->cap_tree) {
            w->cap_tree = w->cap_tree ? 2 * w->cap_tree : 64;
            w->tree = (tree_entry *)realloc(w->tree, sizeof(tree_entry) * (size_t)w->cap_tree);
        }
        tree_entry *e = &w->tree[w->n_tree++];
        memcpy(e->bmin, mn, sizeof mn);
        memcpy(e->bmax, mx, sizeof mx);
        e->is_gate = is_gate;
        e->object_id = id;
        e->obb_index = i;
    }
}

/* World::removeObject, src/World.cpp:69-78 */
static int remove_object(orc_world *w, int id, int is_gate, int n_obbs, double inflate)
{
    for (int i = 0; i < n_obbs; ++i) {
        map_obb *o = map_find(w, is_gate, id, i);
        if (!o) return ORC_ERR_NO_SUCH_OBB; /* obbs.at() throws, :74 */
        double mn[3], mx[3];
        obb_aabb(o->center, o->half, o->rot, o->filling, inflate, mn, mx);
        for (int t = 0; t < w->n_tree; ++t) { /* index.remove(value): one equal value, :75 */
            tree_entry *e = &w->tree[t];
            if (e->is_gate == is_gate && e->object_id == id && e->obb_index == i &&
                e->bmin[0] == mn[0] && e->bmin[1] == mn[1] && e->bmin[2] == mn[2] &&
                e->bmax[0] == mx[0] && e->bmax[1] == mx[1] && e->bmax[2] == mx[2]) {
                w->tree[t] = w->tree[--w->n_tree];
                break;
            }
        }
        *o = w->map[--w->n_map]; /* obbs.erase, :76 */
    }
    return ORC_OK;
}

/* World::addGatePrivateOperation, src/World.cpp:13-35 */
static int add_gate_private(orc_world *w, int gate_id, const double pose7[7], int is_update)
{
    double pos[3] = {pose7[0], pose7[1], 0.0}; /* :15-16 */
    double rot[3] = {pose7[3], pose7[4], pose7[5]};
    int type = (int)pose7[6]; /* :18 */
    if (type < 0 || type >= MAX_GATE_TYPES || w->gate_geom[type].n == 0) return ORC_ERR_NO_GEOMETRY;
    const geom_list *g = &w->gate_geom[type];
    map_obb *objs = (map_obb *)malloc(sizeof(map_obb) * (size_t)g->n);
    int rc = build_object(g, pos, rot, objs); /* :23 */
    if (rc == ORC_OK && is_update) rc = remove_object(w, gate_id, 1, g->n, w->inflate_gate); /* :25-32 */
    if (rc == ORC_OK) add_object(w, objs, g->n, gate_id, 1, w->inflate_gate);                /* :34 */
    free(objs);
    return rc;
}

int orc_world_add_gate(orc_world *w, int gate_id, const double pose7[7])
{
    return add_gate_private(w, gate_id, pose7, 0); /* src/World.cpp:37-40 */
}

int orc_world_update_gate(orc_world *w, int gate_id, const double pose7[7])
{
    return add_gate_private(w, gate_id, pose7, 1); /* src/World.cpp:42-45 */
}

int orc_world_add_obstacle(orc_world *w, int obstacle_id, const double pose6[6]) /* src/World.cpp:47-55 */
{
    double pos[3] = {pose6[0], pose6[1], pose6[2]};
    double rot[3] = {pose6[3], pose6[4], pose6[5]};
    const geom_list *g = &w->obstacle_geom;
    if (g->n == 0) return ORC_ERR_NO_GEOMETRY;
    map_obb *objs = (map_obb *)malloc(sizeof(map_obb) * (size_t)g->n);
    int rc = build_object(g, pos, rot, objs);
    if (rc == ORC_OK) add_object(w, objs, g->n, obstacle_id, 0, w->inflate_obstacle);
    free(objs);
    return rc;
}

int orc_world_parse_gates_and_obstacles(orc_world *w, const double *gates7, int n_gates,
                                        const double *obstacles6, int n_obstacles)
{
    orc_world_reset(w); /* src/PathPlanner.cpp:63 */
    for (int i = 0; i < n_gates; ++i) {
        double g[7];
        memcpy(g, gates7 + 7 * i, sizeof g);
        g[2] = 0.0; /* :68 */
        int rc = orc_world_add_gate(w, i, g);
        if (rc) return rc;
    }
    for (int i = 0; i < n_obstacles; ++i) {
        int rc = orc_world_add_obstacle(w, i, obstacles6 + 6 * i);
        if (rc) return rc;
    }
    return ORC_OK;
}

int orc_world_size(const orc_world *w) { return w->n_tree; }

/* Flatten R-tree value i together with the OBB its key resolves to. */
int orc_world_get(const orc_world *w, int i, orc_obb *out)
{
    if (i < 0 || i >= w->n_tree) return ORC_ERR_NO_SUCH_OBB;
    const tree_entry *e = &w->tree[i];
    const map_obb *o = map_find(w, e->is_gate, e->object_id, e->obb_index);
    if (!o) return ORC_ERR_NO_SUCH_OBB;
    memcpy(out->center, o->center, sizeof o->center);
    memcpy(out->half, o->half, sizeof o->half);
    memcpy(out->rot, o->rot, sizeof o->rot);
    memcpy(out->aabb_min, e->bmin, sizeof e->bmin);
    memcpy(out->aabb_max, e->bmax, sizeof e->bmax);
    out->inflate = e->is_gate ? w->inflate_gate : w->inflate_obstacle;
    out->filling = o->filling;
    out->is_gate = e->is_gate;
    out->object_id = e->object_id;
    out->obb_index = e->obb_index;
    return ORC_OK;
}

/* -------------------------------------------------------------- predicates */

/* Boost within(point, box): strict on every axis (bgi::contains, src/World.cpp:83) */
static inline int within_strict(const tree_entry *e, const double p[3])
{
    return p[0] > e->bmin[0] && p[0] < e->bmax[0] && p[1] > e->bmin[1] && p[1] < e->bmax[1] &&
           p[2] > e->bmin[2] && p[2] < e->bmax[2];
}

/* Boost intersects(box, box): closed boxes (src/World.cpp:143) */
static inline int boxes_intersect(const tree_entry *e, const double mn[3], const double mx[3])
{
    for (int i = 0; i < 3; ++i) {
        if (mx[i] < e->bmin[i]) return 0;
        if (mn[i] > e->bmax[i]) return 0;
    }
    return 1;
}

int orc_check_point(const orc_world *w, const double p[3], int can_pass_gate) /* src/World.cpp:80-104 */
{
    for (int t = 0; t < w->n_tree; ++t) {
        const tree_entry *e = &w->tree[t];
        if (!within_strict(e, p)) continue;                                          /* :83 */
        const map_obb *o = map_find(w, e->is_gate, e->object_id, e->obb_index);      /* :87 */
        double inflate = e->is_gate ? w->inflate_gate : w->inflate_obstacle;         /* :89-90 */
        if (o->filling && can_pass_gate) continue;                                   /* :92-95 */
        if (point_hit(o->center, o->half, o->rot, o->filling, p, inflate)) return 0; /* :97-100 */
    }
    return 1;
}

int orc_check_point_mindist(const orc_world *w, const double p[3], double min_distance) /* :106-128 */
{
    for (int t = 0; t < w->n_tree; ++t) {
        const tree_entry *e = &w->tree[t];
        if (!within_strict(e, p)) continue; /* :109 */
        const map_obb *o = map_find(w, e->is_gate, e->object_id, e->obb_index);
        if (o->filling) continue;                                                         /* :116-119 */
        if (point_hit(o->center, o->half, o->rot, o->filling, p, min_distance)) return 0; /* :121 */
    }
    return 1;
}

int orc_check_ray(const orc_world *w, const double a[3], const double b[3], int can_pass_gate) /* :130-163 */
{
    double mn[3], mx[3];
    for (int i = 0; i < 3; ++i) { /* rowwise().minCoeff()/maxCoeff(), :134-139 */
        mn[i] = a[i];
        mx[i] = a[i];
        if (b[i] < mn[i]) mn[i] = b[i];
        if (b[i] > mx[i]) mx[i] = b[i];
    }
    for (int t = 0; t < w->n_tree; ++t) {
        const tree_entry *e = &w->tree[t];
        if (!boxes_intersect(e, mn, mx)) continue; /* :143 */
        const map_obb *o = map_find(w, e->is_gate, e->object_id, e->obb_index);
        if (o->filling && can_pass_gate) continue;                                      /* :150-153 */
        double inflate = e->is_gate ? w->inflate_gate : w->inflate_obstacle;            /* :155 */
        if (ray_hit(o->center, o->half, o->rot, o->filling, a, b, inflate)) return 0;   /* :157-160 */
    }
    return 1;
}

/* --------------------------------------------------------------- batching */

/* Flattened snapshot so the batched loops do not pay map_find per candidate
 * (set semantics of the R-tree query make the traversal order irrelevant). */
typedef struct {
    int n;
    orc_obb *o;
} flat_world;

static flat_world flatten(const orc_world *w)
{
    flat_world f;
    f.n = w->n_tree;
    f.o = (orc_obb *)malloc(sizeof(orc_obb) * (size_t)(f.n > 0 ? f.n : 1));
    for (int i = 0; i < f.n; ++i) orc_world_get(w, i, &f.o[i]);
    return f;
}

static inline int flat_point(const flat_world *f, const double p[3], int mode, int can_pass_gate,
                             double min_distance)
{
    for (int t = 0; t < f->n; ++t) {
        const orc_obb *o = &f->o[t];
        if (!(p[0] > o->aabb_min[0] && p[0] < o->aabb_max[0] && p[1] > o->aabb_min[1] &&
              p[1] < o->aabb_max[1] && p[2] > o->aabb_min[2] && p[2] < o->aabb_max[2]))
            continue;
        if (mode == 0) {
            if (o->filling && can_pass_gate) continue;
            if (point_hit(o->center, o->half, o->rot, o->filling, p, o->inflate)) return 0;
        } else {
            if (o->filling) continue;
            if (point_hit(o->center, o->half, o->rot, o->filling, p, min_distance)) return 0;
        }
    }
    return 1;
}

static inline int flat_ray(const flat_world *f, const double a[3], const double b[3],
                           int can_pass_gate)
{
    double mn[3], mx[3];
    for (int i = 0; i < 3; ++i) {
        mn[i] = a[i];
        mx[i] = a[i];
        if (b[i] < mn[i]) mn[i] = b[i];
        if (b[i] > mx[i]) mx[i] = b[i];
    }
    for (int t = 0; t < f->n; ++t) {
        const orc_obb *o = &f->o[t];
        int disjoint = 0;
        for (int i = 0; i < 3; ++i)
            if (mx[i] < o->aabb_min[i] || mn[i] > o->aabb_max[i]) disjoint = 1;
        if (disjoint) continue;
        if (o->filling && can_pass_gate) continue;
        if (ray_hit(o->center, o->half, o->rot, o->filling, a, b, o->inflate)) return 0;
    }
    return 1;
}

typedef struct {
    const flat_world *f;
    const double *a, *b;
    size_t word_begin, word_end, n;
    int mode; /* 0 point/canPassGate, 1 point/minDistance, 2 ray */
    int can_pass_gate;
    double min_distance;
    uint32_t *mask;
} job;

static void *job_run(void *arg)
{
    job *j = (job *)arg;
    for (size_t wd = j->word_begin; wd < j->word_end; ++wd) {
        uint32_t bits = 0;
        size_t base = wd * 32;
        for (int k = 0; k < 32 && base + (size_t)k < j->n; ++k) {
            size_t i = base + (size_t)k;
            int ok;
            if (j->mode == 2)
                ok = flat_ray(j->f, j->a + 3 * i, j->b + 3 * i, j->can_pass_gate);
            else
                ok = flat_point(j->f, j->a + 3 * i, j->mode, j->can_pass_gate, j->min_distance);
            bits |= (uint32_t)ok << k;
        }
        j->mask[wd] = bits;
    }
    return NULL;
}

static void run_batch(const orc_world *w, const double *a, const double *b, size_t n, int mode,
                      int can_pass_gate, double min_distance, uint32_t *mask, int threads)
{
    flat_world f = flatten(w);
    size_t words = (n + 31) / 32;
    if (threads < 1) threads = 1;
    if ((size_t)threads > words) threads = words ? (int)words : 1;
    job *jobs = (job *)malloc(sizeof(job) * (size_t)threads);
    pthread_t *tid = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)threads);
    for (int t = 0; t < threads; ++t) {
        job *j = &jobs[t];
        j->f = &f;
        j->a = a;
        j->b = b;
        j->n = n;
        j->word_begin = words * (size_t)t / (size_t)threads;
        j->word_end = words * (size_t)(t + 1) / (size_t)threads;
        j->mode = mode;
        j->can_pass_gate = can_pass_gate;
        j->min_distance = min_distance;
        j->mask = mask;
        if (threads > 1) pthread_create(&tid[t], NULL, job_run, j);
    }
    if (threads > 1)
        for (int t = 0; t < threads; ++t) pthread_join(tid[t], NULL);
    else
        job_run(&jobs[0]);
    free(jobs);
    free(tid);
    free(f.o);
}

void orc_check_states(const orc_world *w, const double *xyz, size_t n, int can_pass_gate,
                      uint32_t *mask, int threads)
{
    run_batch(w, xyz, NULL, n, 0, can_pass_gate, 0.0, mask, threads);
}

void orc_check_states_mindist(const orc_world *w, const double *xyz, size_t n, double min_distance,
                              uint32_t *mask, int threads)
{
    run_batch(w, xyz, NULL, n, 1, 0, min_distance, mask, threads);
}

void orc_check_segments(const orc_world *w, const double *a, const double *b, size_t n,
                        int can_pass_gate, uint32_t *mask, int threads)
{
    run_batch(w, a, b, n, 2, can_pass_gate, 0.0, mask, threads);
}

/* PathPlanner::checkTrajectoryValidity, src/PathPlanner.cpp:267-280 */
int orc_check_trajectory(const orc_world *w, const double *col_major, size_t rows, size_t ld,
                         double min_distance, uint32_t *sample_mask, long *first_invalid)
{
    long first = -1;
    if (sample_mask) memset(sample_mask, 0, sizeof(uint32_t) * ((rows + 31) / 32));
    for (size_t i = 0; i < rows; ++i) {
        double p[3] = {col_major[0 * ld + i], col_major[3 * ld + i], col_major[6 * ld + i]}; /* :272-273 */
        int ok = orc_check_point_mindist(w, p, min_distance);                                 /* :274 */
        if (!ok && first < 0) first = (long)i;
        if (sample_mask)
            sample_mask[i >> 5] |= (uint32_t)ok << (i & 31);
        else if (!ok)
            break;
    }
    if (first_invalid) *first_invalid = first;
    return first < 0;
}

/* ------------------------------------------------- polynomial trajectories */

/* Polynomial::evaluate(t, 0), polynomial.h:136-150.  For derivative 0 the base
 * coefficient row is all ones (polynomial.cpp:145-149) and row[j]*c[j] == c[j]. */
double orc_poly_eval(const double *c, int n, double t)
{
    double result = 1.0 * c[n - 1];
    for (int j = n - 2; j >= 0; --j) {
        result *= t;
        result += 1.0 * c[j];
    }
    return result;
}

double orc_traj_max_time(int n_segments, const double *durations) /* trajectory.h:64-71 */
{
    double max_time = 0.0;
    for (int i = 0; i < n_segments; ++i) max_time += durations[i];
    return max_time;
}

/* Trajectory::evaluateRange, trajectory.cpp:81-141 */
size_t orc_traj_evaluate_range(int n_segments, int n_coeffs, const double *coeffs,
                               const double *durations, double t_start, double t_end, double dt,
                               double *out_xyz, double *out_times, size_t max_out)
{
    size_t count = 0;
    double accumulated_time = 0.0;
    int i = 0;
    for (i = 0; i < n_segments; ++i) { /* :99-110 */
        accumulated_time += durations[i];
        if (accumulated_time > t_start) break;
    }
    if (t_start > accumulated_time) return 0; /* :111-114 */
    if (i >= n_segments) return 0;            /* (reference would read past the end) */
    accumulated_time -= durations[i];         /* :117 */
    double time_in_segment = t_start - accumulated_time;
    while (accumulated_time < t_end) { /* :121-140 */
        if (time_in_segment > durations[i]) {
            time_in_segment = time_in_segment - durations[i];
            i++;
            if (i >= n_segments) break;
            continue;
        }
        if (count < max_out) {
            for (int d = 0; d < 3; ++d) /* Segment::evaluate, segment.cpp:51-58 */
                out_xyz[3 * count + d] =
                    orc_poly_eval(coeffs + ((size_t)i * 3 + (size_t)d) * (size_t)n_coeffs, n_coeffs,
                                  time_in_segment);
            if (out_times) out_times[count] = accumulated_time;
        }
        count++;
        time_in_segment += dt;
        accumulated_time += dt;
    }
    return count;
}

int orc_check_poly_traj(const orc_world *w, int n_segments, int n_coeffs, const double *coeffs,
                        const double *durations, double dt, double min_distance,
                        uint32_t *sample_mask, size_t mask_words, size_t *n_samples_out,
                        long *first_invalid)
{
    double t_end = orc_traj_max_time(n_segments, durations);
    size_t n = orc_traj_evaluate_range(n_segments, n_coeffs, coeffs, durations, 0.0, t_end, dt, NULL,
                                       NULL, 0);
    double *xyz = (double *)malloc(sizeof(double) * 3 * (n ? n : 1));
    orc_traj_evaluate_range(n_segments, n_coeffs, coeffs, durations, 0.0, t_end, dt, xyz, NULL, n);
    if (sample_mask) memset(sample_mask, 0, sizeof(uint32_t) * mask_words);
    long first = -1;
    for (size_t i = 0; i < n; ++i) {
        int ok = orc_check_point_mindist(w, xyz + 3 * i, min_distance);
        if (!ok && first < 0) first = (long)i;
        if (sample_mask && (i >> 5) < mask_words) sample_mask[i >> 5] |= (uint32_t)ok << (i & 31);
    }
    free(xyz);
    if (n_samples_out) *n_samples_out = n;
    if (first_invalid) *first_invalid = first;
    return first < 0;
}

/* ------------------------------------------------------ synthetic inputs */

uint64_t orc_mix64(uint64_t seed, uint64_t counter)
{
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (counter + 1);
    z ^= z >> 30;
    z *= 0xBF58476D1CE4E5B9ull;
    z ^= z >> 27;
    z *= 0x94D049BB133111EBull;
    z ^= z >> 31;
    return z;
}

static inline double u01(uint64_t seed, uint64_t index, unsigned lane)
{
    return (double)(orc_mix64(seed, index * 8 + lane) >> 11) * 0x1.0p-53;
}

void orc_gen_states(uint64_t seed, uint64_t first, size_t n, const double lo[3], const double hi[3],
                    double *xyz)
{
    for (size_t k = 0; k < n; ++k)
        for (unsigned d = 0; d < 3; ++d)
            xyz[3 * k + d] = lo[d] + u01(seed, first + k, d) * (hi[d] - lo[d]);
}

void orc_gen_segments(uint64_t seed, uint64_t first, size_t n, const double lo[3],
                      const double hi[3], double max_len, double *a, double *b)
{
    for (size_t k = 0; k < n; ++k) {
        uint64_t idx = first + k;
        double v[3];
        for (unsigned d = 0; d < 3; ++d) {
            a[3 * k + d] = lo[d] + u01(seed, idx, d) * (hi[d] - lo[d]);
            v[d] = 2.0 * u01(seed, idx, 3 + d) - 1.0;
        }
        double len2 = (v[0] * v[0] + v[1] * v[1]) + v[2] * v[2];
        double norm = sqrt(len2);
        double L = max_len * (1.0 - u01(seed, idx, 6)); /* (0, max_len] */
        for (unsigned d = 0; d < 3; ++d) {
            double u = norm > 0.0 ? v[d] / norm : 0.0;
            double e = a[3 * k + d] + L * u;
            if (e < lo[d]) e = lo[d];
            if (e > hi[d]) e = hi[d];
            b[3 * k + d] = e;
        }
    }
}
