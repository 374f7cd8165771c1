"""Where config 1's pre_compute_traj spends its host time, with and without glibc's mmap threshold raised
(with a CUDA context alive every munmap of a numpy temporary goes through the driver's mmu notifier)."""
import cProfile, ctypes, pstats, sys, time
sys.path.insert(0, ".")
import epp_b200, online_traj_planner as otp
config = epp_b200.load_config(); tr = config["track"]
otp.OnlineTrajGenerator.path_segments_folder = None
gen = otp.OnlineTrajGenerator(tr["start"], tr["goal"], tr["gates"], tr["obstacles"], config)
gen.pre_compute_traj(0.0)
def timed(tag):
    ts = []
    for _ in range(5):
        t = time.perf_counter(); gen.pre_compute_traj(0.0); ts.append(time.perf_counter() - t)
    print(tag, ["%.1f ms" % (1e3 * x) for x in ts])
timed("default malloc")
libc = ctypes.CDLL("libc.so.6")
print("mallopt", libc.mallopt(-3, 1 << 30), libc.mallopt(-1, 1 << 30))   # M_MMAP_THRESHOLD, M_TRIM_THRESHOLD
timed("mmap threshold 1 GiB")
pr = cProfile.Profile(); pr.enable(); gen.pre_compute_traj(0.0); pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(14)
print(gen.pathPlanner.last_stats)
