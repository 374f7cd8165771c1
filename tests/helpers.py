"""Shared helpers for the parity tests: identical worlds on the oracle and the GPU path, and
adversarial inputs around every OBB / stored-AABB face."""
import copy

import numpy as np


def filling_config(config):
    """Config variant with a 'filling' box in each portal opening (the shipped geometry has none;
    the type exists in the code: include/OBB.h:54-57, src/World.cpp:92-95,116-119)."""
    c = copy.deepcopy(config)
    c["component_geometry"]["small_portal"]["opening"] = {
        "position": [0, 0, 0.525], "size": [0.45, 0.05, 0.45], "type": "filling", "name": "opening"}
    c["component_geometry"]["large_portal"]["opening"] = {
        "position": [0, 0, 1.0], "size": [0.45, 0.05, 0.45], "type": "filling", "name": "opening"}
    return c


def build_pair(oracle, epp, config, gates, obstacles, broadphase=0):
    ow = oracle.OracleWorld(config)
    ow.parse_gates_and_obstacles(gates, obstacles)
    gw = epp.World(config, broadphase=broadphase)
    for i, g in enumerate(gates):
        g = list(g)
        g[2] = 0.0
        gw.addGate(i, g)
    for i, o in enumerate(obstacles):
        gw.addObstacle(i, o)
    return ow, gw


def ulp_steps(x, ks=(-2, -1, 0, 1, 2)):
    out = []
    for k in ks:
        v = np.float64(x)
        for _ in range(abs(k)):
            v = np.nextafter(v, np.inf if k > 0 else -np.inf)
        out.append(float(v))
    return out


def adversarial_points(obbs, rng, per_face=6):
    """Points on / 1-2 ulp around every stored-AABB face and every (inflated) OBB face."""
    pts = []
    for o in obbs:
        mn, mx = np.array(o["aabb_min"]), np.array(o["aabb_max"])
        c, h = np.array(o["center"]), np.array(o["half"])
        R = np.array(o["rot"]).reshape(3, 3)
        for axis in range(3):
            for face in (mn[axis], mx[axis]):
                for _ in range(per_face):
                    p = mn + rng.random(3) * (mx - mn)
                    for v in ulp_steps(face):
                        q = p.copy()
                        q[axis] = v
                        pts.append(q)
        for infl in (0.0, o["inflate"], 0.1):
            H = h + (0.0 if o["filling"] else infl)
            for axis in range(3):
                for sgn in (-1.0, 1.0):
                    for _ in range(per_face):
                        l = (rng.random(3) * 2 - 1) * H
                        l[axis] = sgn * H[axis]
                        w = R @ l + c
                        for k in (-1, 0, 1):
                            q = w.copy()
                            q[axis] = ulp_steps(q[axis], (k,))[0]
                            pts.append(q)
    return np.array(pts)


def adversarial_segments(obbs, rng, lo, hi, per_obb=24):
    """Zero-length segments, |dir| straddling 1e-6 in a box frame, segments whose bounding box
    only touches a stored AABB, and segments through / grazing the boxes."""
    A, B = [], []
    for o in obbs:
        mn, mx = np.array(o["aabb_min"]), np.array(o["aabb_max"])
        c = np.array(o["center"])
        R = np.array(o["rot"]).reshape(3, 3)
        ext = mx - mn
        for _ in range(per_obb):
            p = mn - 0.3 + rng.random(3) * (ext + 0.6)
            A.append(p); B.append(p.copy())                                   # zero length
            d = (rng.random(3) * 2 - 1)
            A.append(p); B.append(p + d * rng.random() * 1.5)                 # random through
            for eps in (0.5e-6, 0.999999e-6, 1e-6, 1.000001e-6, 2e-6):        # near-parallel in box frame
                axis = rng.integers(3)
                l = (rng.random(3) * 2 - 1) * 0.8
                l[axis] = eps * (1 if rng.random() < 0.5 else -1)
                A.append(p); B.append(p + R @ l)
            axis = rng.integers(3)                                            # bbox touching the stored AABB
            q0 = mn + rng.random(3) * ext
            q1 = q0 + (rng.random(3) * 2 - 1) * 0.5
            for v in ulp_steps(mx[axis], (-1, 0, 1)):
                a, b = q0.copy(), q1.copy()
                a[axis] = v
                b[axis] = v + rng.random() * 0.5
                A.append(a); B.append(b)
            for v in ulp_steps(mn[axis], (-1, 0, 1)):
                a, b = q0.copy(), q1.copy()
                a[axis] = v
                b[axis] = v - rng.random() * 0.5
                A.append(a); B.append(b)
            A.append(c + (rng.random(3) * 2 - 1) * 2.0); B.append(c + (rng.random(3) * 2 - 1) * 2.0)
    return np.array(A), np.array(B)
