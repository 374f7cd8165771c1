"""The reference's Python surface (src/pybind.cpp:10-27) on the B200 path: module names, class, the six methods with
their keyword names, error behaviour; and, on the GPU, one full pre-compute + gate update on the config.json track
(BASELINE.json configs[0]) with every planner decision re-validated by the CPU oracle."""
import inspect

import numpy as np
import pytest

import epp_b200


def test_module_surface():
    import online_traj_planner as otp
    import polynomial_trajectory as pt
    assert pt.OnlineTrajGenerator is otp.OnlineTrajGenerator
    g = otp.OnlineTrajGenerator
    assert list(inspect.signature(g.__init__).parameters)[1:] == [
        "start", "goal", "nominalGatePositionAndType", "nominalObstaclePosition", "configPath"]
    assert list(inspect.signature(g.pre_compute_traj).parameters)[1:] == ["takeoffTime"]
    assert list(inspect.signature(g.update_gate_pos).parameters)[1:] == [
        "gateId", "newPose", "dronePos", "nextGateWithinRange", "flightTime"]
    assert list(inspect.signature(g.sample_traj).parameters)[1:] == ["currentTime"]
    for name in ("get_traj_end_time", "get_planned_traj"):
        assert callable(getattr(g, name))
    assert otp.Vector3d([1, 2, 3]).shape == (3,) and otp.MatrixXd([[1, 2], [3, 4]]).shape == (2, 2)


@pytest.fixture(scope="module")
def generator(config, tmp_path_factory):
    import online_traj_planner as otp
    t = config["track"]
    otp.OnlineTrajGenerator.path_segments_folder = str(tmp_path_factory.mktemp("dumps") / "path_segments")
    return otp.OnlineTrajGenerator(start=otp.Vector3d(t["start"]), goal=otp.Vector3d(t["goal"]),
                                   nominalGatePositionAndType=otp.MatrixXd(t["gates"]),
                                   nominalObstaclePosition=otp.MatrixXd(t["obstacles"]), configPath=config)


def _passes_gates_in_order(gen, traj):
    pos, row = traj[:, [0, 3, 6]], 0
    for gate_id in range(gen.gates.shape[0]):
        while row + 1 < pos.shape[0] and not gen.checkGatePassed(pos[row], pos[row + 1], gate_id):
            row += 1
        if row + 1 >= pos.shape[0]:
            return False
    return True


@pytest.mark.gpu
def test_pre_compute_and_gate_update_on_the_track(oracle, config, generator):
    gen = generator
    with pytest.raises(RuntimeError, match="No trajectory data available."):
        gen.sample_traj(currentTime=0.0)
    with pytest.raises(RuntimeError, match="No trajectory data available."):
        gen.get_traj_end_time()
    gen.pre_compute_traj(takeoffTime=0.5)

    # every edge the planner kept is valid for the reference's predicate (CPU oracle), canPassGate as planned
    ow = oracle.OracleWorld(config)
    ow.parse_gates_and_obstacles(gen.gates, gen.obstacles)
    assert len(gen.pathSegments) == gen.gates.shape[0] + 1
    for seg in gen.pathSegments:
        assert seg.shape[0] >= 2
        ok = ow.check_segments(seg[:-1], seg[1:], False, threads=1)
        assert np.unpackbits(ok.view(np.uint8), bitorder="little")[: seg.shape[0] - 1].all()
    assert np.allclose(gen.pathSegments[0][0], config["track"]["start"])
    assert np.allclose(gen.pathSegments[-1][-1], config["track"]["goal"])

    traj = gen.get_planned_traj()
    assert traj.ndim == 2 and traj.shape[1] == 10 and traj.shape[0] > 20
    assert traj[0, 9] == 0.5 and np.all(np.diff(traj[:, 9]) > 0)
    dt = config["trajectory_generator_properties"]["sampling_interval"]
    assert np.allclose(np.diff(traj[:, 9]), dt)
    assert np.allclose(traj[0, [0, 3, 6]], config["track"]["start"]) and np.allclose(traj[0, [1, 4, 7]], 0.0)
    assert np.linalg.norm(traj[-1, [0, 3, 6]] - np.asarray(config["track"]["goal"])) < 0.05
    assert _passes_gates_in_order(gen, traj)
    assert gen.get_traj_end_time() == traj[-1, 9]
    k = traj.shape[0] // 2
    assert np.array_equal(gen.sample_traj(currentTime=traj[k, 9] + 0.01), traj[k])

    # gate 1 is observed 15 cm from its nominal place just after the drone has flown through gate 0 (the reference
    # re-plans from the drone to gate `gateId`, i.e. it assumes that gate is the next one)
    pos = traj[:, [0, 3, 6]]
    r0 = next(i for i in range(pos.shape[0] - 1) if gen.checkGatePassed(pos[i], pos[i + 1], 0)) + 2
    assert gen.update_gate_pos(gateId=1, newPose=gen.gates[1, :6], dronePos=pos[r0], nextGateWithinRange=False,
                               flightTime=traj[r0, 9]) is False
    new_pose = gen.gates[1, :6].copy()
    new_pose[:2] += [0.12, 0.09]
    changed = gen.update_gate_pos(gateId=1, newPose=new_pose, dronePos=pos[r0], nextGateWithinRange=True,
                                  flightTime=traj[r0, 9])
    assert changed is True
    assert np.allclose(gen.gates[1, :2], new_pose[:2])
    new_traj = gen.get_planned_traj()
    assert np.array_equal(new_traj[: r0 + 1], traj[: r0 + 1])      # the flown part is kept
    assert np.all(np.diff(new_traj[:, 9]) > -1e-9)
    assert _passes_gates_in_order(gen, new_traj)                   # gate 0 in the kept part, then the MOVED gate 1, 2, 3
    assert np.linalg.norm(new_traj[-1, [0, 3, 6]] - np.asarray(config["track"]["goal"])) < 0.05
    # the PathWriter dumps of the run (src/OnlineTrajGenerator.cpp:23,29,47,90,147,277,369)
    import os
    folder = gen.pathWriter.folderPath
    assert sorted(os.listdir(folder)) == ["checkpoints.txt", "gates.txt", "obstacles.txt", "path_0.txt", "path_1.txt"]
    assert len(open(os.path.join(folder, "gates.txt")).readlines()) == gen.gates.shape[0] + 1
    assert len(open(os.path.join(folder, "checkpoints.txt")).readlines()) == 2 * gen.gates.shape[0] + 2
    assert np.allclose(np.loadtxt(os.path.join(folder, "path_0.txt")), gen.prunedPath, atol=1e-4, rtol=1e-5)
    # the same observation again is ignored (src/OnlineTrajGenerator.cpp:131-134)
    assert gen.update_gate_pos(gateId=1, newPose=new_pose, dronePos=pos[r0], nextGateWithinRange=True,
                               flightTime=traj[r0, 9]) is False


@pytest.mark.gpu
def test_prune_waypoints_equals_the_reference_loop(oracle, config):
    """PathPlanner.pruneWaypoints answers its ray checks from one batch; the result equals the reference's loop
    (src/PathPlanner.cpp:232-265) run one checkRayValid(…, true) at a time on the CPU oracle."""
    import online_traj_planner as otp
    t = config["track"]
    pp = otp.PathPlanner(t["gates"], t["obstacles"], config)
    ow = oracle.OracleWorld(config)
    ow.parse_gates_and_obstacles(np.asarray(t["gates"], float), np.asarray(t["obstacles"], float))
    rng = np.random.default_rng(8)
    for _ in range(10):
        w = np.cumsum(rng.normal(size=(12, 3)) * 0.25, axis=0) + [0.0, 0.0, 1.0]
        want, ref, idx = [w[0]], 0, 2
        while idx < len(w):
            ok = ow.check_segments(w[ref][None], w[idx][None], True, threads=1)[0] & 1
            if not ok:
                want.append(w[idx - 1])
                ref = idx - 1
            idx += 1
        want.append(w[-1])
        assert np.array_equal(pp.pruneWaypoints(w), np.array(want))


@pytest.mark.gpu
def test_call_trace_of_one_solve_replays_bit_for_bit(oracle, config, tmp_path):
    """record EVERY question one pre_compute_traj asks the world — the states and edges the
    planner got "invalid" for as well as the ones it kept (src/PathPlanner.cpp:80-158 asks them through
    StateValidator::isValid / MotionValidator::checkMotion, src/StateValidator.cpp:7-13, src/MotionValidator.cpp:8-17) —
    and replay the trace through the CPU oracle, the reference's own code where it is built, and, one call at a time,
    through the single-unit entry points the OMPL virtuals use.  Masks must be equal everywhere."""
    import time
    import online_traj_planner as otp
    t = config["track"]
    otp.OnlineTrajGenerator.path_segments_folder = str(tmp_path / "path_segments")
    gen = otp.OnlineTrajGenerator(t["start"], t["goal"], t["gates"], t["obstacles"], config)
    world = gen.pathPlanner.worldPtr
    trace = []
    real_states, real_segments = world.check_states, world.check_segments

    def rec_states(xyz, can_pass_gate=False):
        m = real_states(xyz, can_pass_gate)
        trace.append(("states", np.array(xyz, dtype=np.float64, copy=True), None, bool(can_pass_gate), np.array(m, copy=True)))
        return m

    def rec_segments(a, b, can_pass_gate=False):
        m = real_segments(a, b, can_pass_gate)
        trace.append(("segments", np.array(a, dtype=np.float64, copy=True), np.array(b, dtype=np.float64, copy=True),
                      bool(can_pass_gate), np.array(m, copy=True)))
        return m

    world.check_states, world.check_segments = rec_states, rec_segments
    try:
        gen.pre_compute_traj(0.0)
    finally:
        world.check_states, world.check_segments = real_states, real_segments
    n_states = sum(q[1].shape[0] for q in trace if q[0] == "states")
    n_edges = sum(q[1].shape[0] for q in trace if q[0] == "segments")
    assert n_states > 1000 and n_edges > 1000

    checkers = [oracle.OracleWorld(config)]
    try:
        import ref_py
        if ref_py.available():
            checkers.append(ref_py.RefWorld(config))
    except ImportError:
        pass
    invalid_answers = 0
    for cw in checkers:
        cw.parse_gates_and_obstacles(gen.gates, gen.obstacles)
        for kind, a, b, cpg, got in trace:
            want = cw.check_states(a, cpg, threads=4) if kind == "states" else cw.check_segments(a, b, cpg, threads=4)
            assert np.array_equal(np.asarray(got, np.uint32), want), (kind, a.shape, cpg)
            invalid_answers += a.shape[0] - int(np.unpackbits(want.view(np.uint8)).sum())
    assert invalid_answers > 0  # the trace holds rejected questions too, not only the kept ones

    # one call at a time through epp_is_valid / epp_check_motion (what OMPL's per-call loop would do): a sample of
    # the trace, answers equal to the batch's bits; per-call latency printed for the record
    rng = np.random.default_rng(5)
    t0, calls = time.perf_counter(), 0
    for kind, a, b, cpg, got in trace:
        bits = np.unpackbits(np.asarray(got, np.uint32).view(np.uint8), bitorder="little")
        for i in rng.choice(a.shape[0], size=min(40, a.shape[0]), replace=False):
            one = world.checkPointValidity(a[i], cpg) if kind == "states" else world.checkRayValid(a[i], b[i], cpg)
            assert one == bool(bits[i]), (kind, i)
            calls += 1
    print("call trace: %d states + %d edges in %d batches; %d single calls replayed, %.1f us per call (incl. binding)"
          % (n_states, n_edges, len(trace), calls, (time.perf_counter() - t0) / calls * 1e6))
