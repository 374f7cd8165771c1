import numpy as np, time, os, sys
from scipy.sparse import coo_matrix
from scipy.sparse.csgraph import dijkstra
from scipy.spatial import cKDTree
print("cpus", len(os.sched_getaffinity(0)), os.cpu_count(), open("/proc/loadavg").read().strip())
rng=np.random.default_rng(0)
pts=rng.uniform([-3,-3,0],[3,3,2],size=(929,3))
def probe(tag):
    for _ in range(2):
        t=time.perf_counter(); pairs=cKDTree(pts).query_pairs(1.1,output_type="ndarray"); t1=time.perf_counter()
        w=np.linalg.norm(pts[pairs[:,0]]-pts[pairs[:,1]],axis=1); t2=time.perf_counter()
        g=coo_matrix((w,(pairs[:,0],pairs[:,1])),shape=(929,929))
        d,p=dijkstra(g,directed=False,indices=0,return_predecessors=True); t3=time.perf_counter()
    print(tag, pairs.shape, "kdtree %.1f ms norm %.1f ms dijkstra %.1f ms" % (1e3*(t1-t), 1e3*(t2-t1), 1e3*(t3-t2)))
probe("before cuda")
sys.path.insert(0, ".")
import epp_b200
probe("after import epp_b200")
config = epp_b200.load_config(); tr = config["track"]
w = epp_b200.World(config); w.parse_gates_and_obstacles(np.asarray(tr["gates"],float), np.asarray(tr["obstacles"],float))
w.check_states(pts, False)
probe("after cuda init")
import online_traj_planner as otp
otp.OnlineTrajGenerator.path_segments_folder = None
gen = otp.OnlineTrajGenerator(tr["start"], tr["goal"], tr["gates"], tr["obstacles"], config)
gen.pre_compute_traj(0.0)
pp = gen.pathPlanner
import time
for i in range(3):
    t=time.perf_counter(); p = pp.planPath(np.asarray(tr["start"],float), np.asarray(tr["gates"][0][:3],float)+[0,0,0.0], 1.0); print("planPath %.1f ms" % (1e3*(time.perf_counter()-t)), pp.last_stats)
orig = pp.planPath
def wrapped(s, g, tl):
    t = time.perf_counter(); r = orig(s, g, tl)
    print("planPath %.1f ms" % (1e3 * (time.perf_counter() - t)), np.round(s, 2), np.round(g, 2), pp.last_stats, None if r is None else len(r))
    return r
pp.planPath = wrapped
gen.pre_compute_traj(0.0)
