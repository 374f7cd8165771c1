"""online_traj_planner — the reference's Python surface (src/pybind.cpp:10-27) on the B200 path.

``OnlineTrajGenerator`` keeps the constructor and the six methods of the pybind class with their keyword names
(``pre_compute_traj``, ``update_gate_pos``, ``sample_traj``, ``get_traj_end_time``, ``get_planned_traj``) and follows
src/OnlineTrajGenerator.cpp step by step; what it calls underneath is this repository's GPU path:

  * world queries           -> epp_b200.World (hand-written sm_100a kernels behind the C ABI; no CPU fallback)
  * PathPlanner::planPath   -> BatchPlanner below: the per-call OMPL loop (src/PathPlanner.cpp:80-158) is replaced by a
                               batch roadmap — one isValid batch over all samples, one checkMotion batch over all
                               candidate edges, shortest path, one checkMotion batch over all shortcuts
  * pruneWaypoints          -> the reference's loop (src/PathPlanner.cpp:232-265) over ONE batch of ray checks
  * generateTrajectory      -> epp_b200.segment_times + epp_b200.minsnap (batched min-snap solve) +
                               epp_b200.sample_trajectories (the three evaluateRange passes), type "snap" only
  * checkTrajectoryValidity -> World.checkTrajectoryValidity (K3a)

Not built: trajectory types "spline" (Eigen::Spline fitting) and "optimal" (external/time_parametrization), the
PathWriter text dumps, the detached re-planning thread (``recalculate_online``; re-planning runs on the caller's thread).
"""
import math

import numpy as np

from . import World, load_config, minsnap, sample_trajectories, segment_times


class Vector3d(np.ndarray):
    """pybind's Vector3d wrapper (src/pybind.cpp:22-23): built from a numpy array."""

    def __new__(cls, a):
        return np.asarray(a, dtype=np.float64).reshape(3).view(cls)


class MatrixXd(np.ndarray):
    """pybind's MatrixXd wrapper (src/pybind.cpp:25-26)."""

    def __new__(cls, a):
        return np.atleast_2d(np.asarray(a, dtype=np.float64)).view(cls)


def _bits(mask, n):
    return np.unpackbits(np.ascontiguousarray(mask).view(np.uint8), bitorder="little")[:n].astype(bool)


class BatchPlanner:
    """Batch roadmap planner over a World (the Python twin of host/src/BatchPlanner.cpp).  Deterministic: samples come
    from the counter-based generator of epp_b200.worlds."""

    def __init__(self, world, lo, hi, can_pass_gate=False, planner_range=0.5, initial_samples=2048,
                 max_samples=1 << 17, seed=0x5EED0001):
        self.world, self.lo, self.hi = world, np.asarray(lo, float), np.asarray(hi, float)
        self.cpg, self.range = bool(can_pass_gate), float(planner_range)
        self.initial_samples, self.max_samples, self.seed = int(initial_samples), int(max_samples), int(seed)
        self.stats = {}

    def solve(self, start, goal, time_limit=2.0):
        import time
        from scipy.sparse import coo_matrix
        from scipy.sparse.csgraph import dijkstra
        from scipy.spatial import cKDTree
        from .worlds import gen_states
        t0 = time.perf_counter()
        start, goal = np.asarray(start, float).reshape(3), np.asarray(goal, float).reshape(3)
        st = dict(rounds=0, samples=0, valid_samples=0, edges_checked=0, valid_edges=0, shortcut_checks=0)
        self.stats = st
        ok = _bits(self.world.check_states(np.stack([start, goal]), self.cpg), 2)
        if not ok.all():
            return None
        volume = float(np.prod(self.hi - self.lo))
        n = self.initial_samples
        while True:
            st["rounds"] += 1
            xyz = np.vstack([start, goal, gen_states(self.seed + st["rounds"], 0, n, self.lo, self.hi),
                             self._terminal_samples(start, goal, n, st["rounds"])])
            valid = _bits(self.world.check_states(xyz, self.cpg), xyz.shape[0])  # one isValid batch
            pts = xyz[valid]
            m = pts.shape[0]
            st["samples"] += n
            st["valid_samples"] += m
            gamma = 2.0 * ((1.0 + 1.0 / 3.0) * volume / (4.0 / 3.0 * math.pi)) ** (1.0 / 3.0)
            r = max(self.range * (2048.0 / n) ** (1.0 / 3.0), gamma * (math.log(m) / m) ** (1.0 / 3.0))
            pairs = cKDTree(pts).query_pairs(r, output_type="ndarray")
            if pairs.shape[0]:
                emask = _bits(self.world.check_segments(pts[pairs[:, 0]], pts[pairs[:, 1]], self.cpg), pairs.shape[0])
            else:
                emask = np.zeros(0, dtype=bool)                                  # one checkMotion batch
            st["edges_checked"] += pairs.shape[0]
            good = pairs[emask]
            st["valid_edges"] += good.shape[0]
            w = np.linalg.norm(pts[good[:, 0]] - pts[good[:, 1]], axis=1)
            graph = coo_matrix((w, (good[:, 0], good[:, 1])), shape=(m, m))
            dist, pred = dijkstra(graph, directed=False, indices=0, return_predecessors=True)
            if np.isfinite(dist[1]):
                chain, v = [], 1
                while v >= 0:
                    chain.append(v)
                    v = pred[v]
                return self.reduce_vertices(pts[chain[::-1]])
            if time.perf_counter() - t0 > time_limit or 2 * n > self.max_samples:
                return None
            n *= 2

    def _terminal_samples(self, start, goal, n, rnd):
        """n/8 extra samples in boxes of half-width range and range/4 around start and goal (generator indices
        n, n+1, ...): checkpoints sit 0.21 m behind a gate, often in a slot between the gate and an obstacle that
        uniform samples only reach after several doublings."""
        from .worlds import gen_states
        nt = max(64, n // 8) & ~3
        j = np.arange(nt)
        c = np.where((j & 1)[:, None] == 0, start[None], goal[None])
        half = np.where((j & 2) == 0, self.range, 0.25 * self.range)[:, None]
        u = gen_states(self.seed + rnd, n, nt, np.zeros(3), np.ones(3))
        return np.clip(c + (2.0 * u - 1.0) * half, self.lo, self.hi)

    def reduce_vertices(self, path):
        """Every shortcut (i, j), j > i + 1, in one edge batch, then the shortest chain over valid shortcuts."""
        k = path.shape[0]
        if k < 3:
            return path
        ii, jj = np.triu_indices(k, 2)
        ok = _bits(self.world.check_segments(path[ii], path[jj], self.cpg), ii.size)
        self.stats["shortcut_checks"] += ii.size
        cost = np.full(k, np.inf)
        parent = np.full(k, -1)
        cost[0] = 0.0
        into = [[j - 1] for j in range(k)]
        for i, j in zip(ii[ok], jj[ok]):
            into[j].append(i)
        for j in range(1, k):
            for i in into[j]:
                c = cost[i] + np.linalg.norm(path[i] - path[j])
                if c < cost[j]:
                    cost[j], parent[j] = c, i
        out, v = [], k - 1
        while v >= 0:
            out.append(path[v])
            v = parent[v]
        return np.array(out[::-1])


class PathPlanner:
    """The pieces of the reference PathPlanner (include/PathPlanner.h) the generator uses."""

    def __init__(self, gates, obstacles, config):
        self.config = config
        self.pp = config["path_planner_properties"]
        wp = config["world_properties"]
        self.lo, self.hi = np.asarray(wp["lower_bound"], float), np.asarray(wp["upper_bound"], float)
        self.worldPtr = World(config)
        self.parseGatesAndObstacles(gates, obstacles)

    def parseGatesAndObstacles(self, gates, obstacles):       # src/PathPlanner.cpp:56-78
        self.worldPtr.parse_gates_and_obstacles(np.asarray(gates, float), np.asarray(obstacles, float))

    def planPath(self, start, goal, timeLimit):               # src/PathPlanner.cpp:80-158
        planner = BatchPlanner(self.worldPtr, self.lo, self.hi, can_pass_gate=self.pp.get("can_pass_gate", False),
                               planner_range=self.pp.get("range", 0.5),
                               initial_samples=max(256, int(self.pp.get("samples_fmt", 1000))))
        path = planner.solve(start, goal, timeLimit)
        self.last_stats = planner.stats
        return path

    def pruneWaypoints(self, waypoints):                      # src/PathPlanner.cpp:232-265
        w = np.asarray(waypoints, float)
        k = w.shape[0]
        if k < 3:
            return w
        # the loop asks checkRayValid(reference, current, true) for pairs (i, j), j >= i + 2: all of them in one batch
        ii, jj = np.triu_indices(k, 2)
        table = np.zeros((k, k), dtype=bool)
        table[ii, jj] = _bits(self.worldPtr.check_segments(w[ii], w[jj], True), ii.size)
        pruned, ref, idx = [w[0]], 0, 2
        while idx < k:
            if not table[ref, idx]:
                pruned.append(w[idx - 1])
                ref = idx - 1
            idx += 1
        pruned.append(w[-1])
        return np.array(pruned)

    def includeGates2(self, segments):                        # src/PathPlanner.cpp:176-230
        segs = [np.asarray(s, float) for s in segments]
        centers = [(segs[i][-1] + segs[i + 1][0]) / 2 for i in range(len(segs) - 1)]
        for i, c in enumerate(centers):
            segs[i] = np.vstack([segs[i], c])
            segs[i + 1] = np.vstack([c, segs[i + 1]])
        method = self.pp.get("path_simplification", "custom")
        flat = []
        for seg in segs:
            if method == "none":
                pruned = seg
            elif method == "custom":
                pruned = self.pruneWaypoints(seg)
            else:
                raise RuntimeError("Unknown pruning method")
            for p in pruned:
                if flat and np.linalg.norm(flat[-1] - p) < 0.05:
                    continue
                flat.append(p)
        return np.array(flat)

    def checkTrajectoryValidity(self, trajectory, minDistance):   # src/PathPlanner.cpp:267-280
        t = np.asarray(trajectory, float)
        if t.shape[0] == 0:
            return True
        return self.worldPtr.checkTrajectoryValidity(t, minDistance)


def generate_trajectory(waypoints, v_max, a_max, sampling_interval, start_time_offset, initial_vel, initial_acc):
    """poly_traj::generateTrajectory (external/poly_traj/src/trajectory_generator.cpp:12-100) on the GPU."""
    wp = np.asarray(waypoints, float)
    if wp.shape[0] < 2:
        raise ValueError("At least two waypoints are required")
    K = wp.shape[0] - 1
    if K > 128:
        raise RuntimeError("generate_trajectory: more than 128 segments (the batched solve holds 128)")
    T = segment_times(wp, v_max, a_max)
    va = np.concatenate([np.asarray(initial_vel, float), np.asarray(initial_acc, float)])[None]
    coeffs = minsnap(wp[None], T[None], va)
    rows = int(T.sum() / sampling_interval) + 8
    return sample_trajectories(coeffs, T[None], sampling_interval, time_offsets=start_time_offset,
                               max_rows=(rows + 31) // 32 * 32)[0]


def _cxx_double(x):
    """operator<<(ostream&, double) with the stream defaults: %g at precision 6."""
    return "%g" % float(x)


class PathWriter:
    """The reference's debug dumps for its plotting scripts (src/PathWriter.cpp:7-113): ``path_<k>.txt`` and
    ``checkpoints.txt`` hold one "x y z " line per point, ``gates.txt`` / ``obstacles.txt`` grow by one
    "id: <i> info: <row>" line per update with the row in Eigen's default stream format (every coefficient
    right-aligned to the widest one, one space between).  As in the reference, an existing folder is emptied of its
    regular files on construction.  ``folderPath=None`` writes nothing."""

    def __init__(self, folderPath):
        import os
        self.folderPath, self.write_count = folderPath, 0
        if folderPath is None:
            return
        if not os.path.exists(folderPath):
            os.makedirs(folderPath)
            print("Folder created: " + folderPath)
        else:
            for name in os.listdir(folderPath):
                full = os.path.join(folderPath, name)
                if os.path.isfile(full) and not os.path.islink(full):
                    os.remove(full)

    def _points(self, fileName, points):
        if self.folderPath is None:
            return
        import os
        with open(os.path.join(self.folderPath, fileName), "w") as f:
            for pt in points:
                f.write("".join(_cxx_double(pt[k]) + " " for k in range(3)) + "\n")

    def _row(self, fileName, ident, row):
        if self.folderPath is None:
            return
        import os
        cells = [_cxx_double(v) for v in np.asarray(row, float).ravel()]
        width = max(len(c) for c in cells) if cells else 0
        with open(os.path.join(self.folderPath, fileName), "a") as f:
            f.write("id: %d info: %s\n" % (ident, " ".join(c.rjust(width) for c in cells)))

    def writePath(self, path):
        self._points("path_%d.txt" % self.write_count, path)
        self.write_count += 1

    def updateGatePos(self, gateId, gateInfo):
        self._row("gates.txt", gateId, gateInfo)

    def updateObstaclePos(self, obstacleId, pose):
        self._row("obstacles.txt", obstacleId, pose)

    def writeCheckpoints(self, checkpoints):
        self._points("checkpoints.txt", checkpoints)


class OnlineTrajGenerator:
    #: folder of the PathWriter dumps (include/OnlineTrajGenerator.h:132); set to None before constructing a
    #: generator to write nothing
    path_segments_folder = "path_segments"

    def __init__(self, start, goal, nominalGatePositionAndType, nominalObstaclePosition, configPath):
        self.config = load_config(configPath) if isinstance(configPath, str) else configPath
        self.pp = self.config["path_planner_properties"]
        self.tg = self.config["trajectory_generator_properties"]
        self.gates = np.array(nominalGatePositionAndType, dtype=np.float64).reshape(-1, 7)
        self.obstacles = np.array(nominalObstaclePosition, dtype=np.float64).reshape(-1, 6)
        self.pathPlanner = PathPlanner(self.gates, self.obstacles, self.config)
        self.pathWriter = PathWriter(self.path_segments_folder)
        for i, gate in enumerate(self.gates):                                 # src/OnlineTrajGenerator.cpp:20-30
            self.pathWriter.updateGatePos(i, gate)
        for i, obstacle in enumerate(self.obstacles):
            self.pathWriter.updateObstaclePos(i, obstacle)
        self.checkpoints = [np.asarray(start, float).reshape(3)]
        for gate in self.gates:                                               # src/OnlineTrajGenerator.cpp:32-46
            center, normal = self.getGateCenterAndNormal(gate)
            off = self.pp["checkpoint_gate_offset"]
            self.checkpoints += [center - off * normal, center + off * normal]
        self.checkpoints.append(np.asarray(goal, float).reshape(3))
        self.pathWriter.writeCheckpoints(self.checkpoints)
        self.pathSegments = []
        self.plannedTraj = np.zeros((0, 10))
        self.gatesObservedWithinRange = set()
        self.trajectoryCurrentlyUpdating = False

    # -- helpers -------------------------------------------------------------------------------------------------
    def getGateCenterAndNormal(self, gate):                                   # src/OnlineTrajGenerator.cpp:49-69
        names = self.config["gate_id_to_name_mapping"]
        height = self.config["component_properties"][names[str(int(gate[6]))]]["height"]
        center = np.array([gate[0], gate[1], gate[2] + height])
        normal = np.array([-math.sin(gate[5]), math.cos(gate[5]), 0.0])
        return center, normal / np.linalg.norm(normal)

    def checkGatePassed(self, pos1, pos2, gateId):                            # src/OnlineTrajGenerator.cpp:229-258
        gate = self.gates[gateId]
        center, _ = self.getGateCenterAndNormal(gate)
        c, s = math.cos(gate[5]), math.sin(gate[5])

        def to_gate(p):
            t = p - center
            return np.array([c * t[0] - s * t[1], s * t[0] + c * t[1], t[2]])
        g1, g2 = to_gate(pos1), to_gate(pos2)
        if g1[1] < 0 and g2[1] > 0:
            mid = (g1 + g2) / 2
            return abs(mid[0]) <= 0.425 and abs(mid[2]) <= 0.425
        return False

    def _generate(self, waypoints, offset, vel, acc):
        if self.tg.get("type", "snap") != "snap":
            raise RuntimeError("Trajectory type not supported")
        return generate_trajectory(waypoints, self.tg["max_velocity"], self.tg["max_acceleration"],
                                   self.tg["sampling_interval"], offset, vel, acc)

    # -- the six pybind methods ----------------------------------------------------------------------------------
    def pre_compute_traj(self, takeoffTime):                                  # src/OnlineTrajGenerator.cpp:71-121
        self.pathSegments = []
        for i in range(0, len(self.checkpoints) - 1, 2):
            path = self.pathPlanner.planPath(self.checkpoints[i], self.checkpoints[i + 1], self.pp["time_limit_offline"])
            if path is None:
                raise RuntimeError("Path not found")
            self.pathSegments.append(path)
        self.prunedPath = self.pathPlanner.includeGates2(self.pathSegments)
        self.pathWriter.writePath(self.prunedPath)
        self.plannedTraj = self._generate(self.prunedPath, takeoffTime, np.zeros(3), np.zeros(3))

    def update_gate_pos(self, gateId, newPose, dronePos, nextGateWithinRange, flightTime):   # :123-226
        if not nextGateWithinRange or gateId in self.gatesObservedWithinRange:
            return False
        dronePos = np.asarray(dronePos, float).reshape(3)
        if not self.pathPlanner.worldPtr.checkPointValidity(dronePos, False):
            return False
        self.gatesObservedWithinRange.add(gateId)
        self.gates[gateId, :6] = np.asarray(newPose, float).reshape(6)
        self.pathPlanner.parseGatesAndObstacles(self.gates, self.obstacles)
        self.pathWriter.updateGatePos(gateId, self.gates[gateId])             # :147
        traj = self.plannedTraj
        startIdx = int(np.argmin(np.abs(traj[:, -1] - flightTime)))
        cpNext = min(2 * gateId + 3, len(self.checkpoints) - 1)
        pos = traj[:, [0, 3, 6]]
        endIdx = int(np.argmin(np.linalg.norm(pos - self.checkpoints[cpNext], axis=1)))
        look = traj[startIdx:max(endIdx, startIdx)]
        passing = any(self.checkGatePassed(look[i, [0, 3, 6]], look[i + 1, [0, 3, 6]], gateId)
                      for i in range(look.shape[0] - 1))
        valid = False
        if passing:
            valid = self.pathPlanner.checkTrajectoryValidity(look, self.pp["min_dist_check_traj_collision"])
        if valid and passing:
            return False
        if self.trajectoryCurrentlyUpdating:
            raise RuntimeError("Call to update trajectory, while previous update is still going on")
        self.trajectoryCurrentlyUpdating = True
        try:
            self.recomputeTraj(gateId, newPose, dronePos, flightTime)
        finally:
            self.trajectoryCurrentlyUpdating = False
        return True

    def recomputeTraj(self, gateId, newPose, dronePos, flightTime):           # src/OnlineTrajGenerator.cpp:260-421
        pre, post, nxt = 2 * gateId + 1, 2 * gateId + 2, 2 * gateId + 3
        center, normal = self.getGateCenterAndNormal(self.gates[gateId])
        off = self.pp["checkpoint_gate_offset"]
        self.checkpoints[pre], self.checkpoints[post] = center - off * normal, center + off * normal
        self.pathWriter.writeCheckpoints(self.checkpoints)                    # :277
        advanced = flightTime
        if self.pp.get("advance_for_calculation", True):
            advanced += self.pp["time_limit_online"] + 0.01
        traj = self.plannedTraj
        later = np.nonzero(traj[:, -1] > advanced)[0]
        idx = int(later[0]) if later.size else traj.shape[0]
        if idx + 1 >= traj.shape[0]:
            return
        part_pre, state = traj[:idx], traj[idx + 1]
        p, v, a = state[[0, 3, 6]], state[[1, 4, 7]], state[[2, 5, 8]]
        if not self.pathPlanner.worldPtr.checkPointValidity(p, bool(self.pp.get("can_pass_gate", False))):
            return   # "Advanced trajectory does not end at valid position. No recomputation and hope for best"
        pre_path = self.pathPlanner.planPath(p, self.checkpoints[pre], self.pp["time_limit_online"])
        if pre_path is None:
            raise RuntimeError("Pre path not found. Exiting")
        post_path = self.pathPlanner.planPath(self.checkpoints[post], self.checkpoints[nxt], self.pp["time_limit_online"])
        if post_path is None:
            raise RuntimeError("Post segment path not found. Exiting")
        self.pathSegments[gateId], self.pathSegments[gateId + 1] = pre_path, post_path
        filled = self.pathPlanner.includeGates2(self.pathSegments[gateId:])
        self.pathWriter.writePath(filled)                                     # :369
        part_post = self._generate(filled, advanced, v, a)
        self.plannedTraj = np.vstack([part_pre, part_post])

    def sample_traj(self, currentTime):                                       # src/OnlineTrajGenerator.cpp:423-439
        if self.plannedTraj.shape[0] == 0:
            raise RuntimeError("No trajectory data available.")
        return self.plannedTraj[int(np.argmin(np.abs(self.plannedTraj[:, -1] - currentTime)))].copy()

    def get_traj_end_time(self):                                              # :441-450
        if self.plannedTraj.shape[0] == 0:
            raise RuntimeError("No trajectory data available.")
        return float(self.plannedTraj[-1, -1])

    def get_planned_traj(self):                                               # :452-460
        if self.plannedTraj.shape[0] == 0:
            raise RuntimeError("No trajectory data available.")
        return self.plannedTraj.copy()

    # C++ names
    preComputeTraj, updateGatePos, sampleTraj = pre_compute_traj, update_gate_pos, sample_traj
    getTrajEndTime, getPlannedTraj = get_traj_end_time, get_planned_traj
