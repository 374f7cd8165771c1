"""Host logic of the batch roadmap planner without a GPU: online_traj_planner.BatchPlanner only
needs an object with check_states / check_segments, so here the CPU oracle answers (tests are the one place allowed to
use it).  Covers the queries of the config.json track, including the one between two checkpoints wedged between a gate
and an obstacle that uniform sampling alone only connected after 7-8 doublings."""
import numpy as np
import pytest

import epp_b200
from epp_b200.online_traj_planner import BatchPlanner, _bits


class OracleBackedWorld:
    def __init__(self, ow):
        self.ow, self.state_batches, self.edge_batches = ow, 0, 0

    def check_states(self, xyz, can_pass_gate=False):
        self.state_batches += 1
        return self.ow.check_states(np.ascontiguousarray(xyz, dtype=np.float64), can_pass_gate, threads=4)

    def check_segments(self, a, b, can_pass_gate=False):
        self.edge_batches += 1
        return self.ow.check_segments(np.ascontiguousarray(a, dtype=np.float64), np.ascontiguousarray(b, dtype=np.float64),
                                      can_pass_gate, threads=4)


@pytest.fixture(scope="module")
def track_planner(oracle, config):
    tr = config["track"]
    ow = oracle.OracleWorld(config)
    ow.parse_gates_and_obstacles(np.asarray(tr["gates"], float), np.asarray(tr["obstacles"], float))
    wp = config["world_properties"]
    world = OracleBackedWorld(ow)
    planner = BatchPlanner(world, wp["lower_bound"], wp["upper_bound"], can_pass_gate=False,
                           planner_range=config["path_planner_properties"].get("range", 0.5), initial_samples=1000)
    return ow, world, planner


QUERIES = [([1.0, 1.0, 0.05], [0.6, -0.85, 0.52]),         # start -> before gate 0
           ([0.0, 0.71, 0.52], [-0.5, -0.29, 1.0]),         # after gate 2 -> before gate 3: both ends in a slot
           ([-0.5, -0.71, 1.0], [0.0, -2.0, 0.5])]          # after gate 3 -> goal


@pytest.mark.parametrize("start,goal", QUERIES)
def test_paths_are_valid_and_found_in_few_rounds(track_planner, start, goal):
    ow, world, planner = track_planner
    path = planner.solve(np.array(start), np.array(goal), time_limit=20.0)
    assert path is not None and path.shape[0] >= 2
    assert np.array_equal(path[0], start) and np.array_equal(path[-1], goal)
    assert _bits(ow.check_states(path, False, threads=1), path.shape[0]).all()
    assert _bits(ow.check_segments(path[:-1], path[1:], False, threads=1), path.shape[0] - 1).all()
    st = planner.stats
    assert st["rounds"] <= 3, st                      # terminal samples: the slot query took 8 rounds without them
    assert st["samples"] == sum(1000 << r for r in range(st["rounds"]))
    # one isValid batch for start/goal + one per round; one checkMotion batch per round + one for the shortcuts
    assert st["edges_checked"] > 0 and st["valid_edges"] <= st["edges_checked"]


def test_planner_is_deterministic_and_batches_its_questions(track_planner):
    ow, world, planner = track_planner
    s, g = np.array(QUERIES[0][0]), np.array(QUERIES[0][1])
    world.state_batches = world.edge_batches = 0
    p1 = planner.solve(s, g, 20.0)
    rounds = planner.stats["rounds"]
    assert world.state_batches == 1 + rounds
    assert world.edge_batches in (rounds, rounds + 1)          # + the shortcut batch when the raw path has > 2 vertices
    p2 = planner.solve(s, g, 20.0)
    assert np.array_equal(p1, p2)


def test_invalid_terminal_fails_without_sampling(track_planner):
    ow, world, planner = track_planner
    gate0 = np.array([0.45, -1.0, 0.4])                # inside the support of gate 0
    assert not _bits(ow.check_states(gate0[None], False, threads=1), 1)[0]
    assert planner.solve(gate0, np.array([0.0, -2.0, 0.5]), 5.0) is None
    assert planner.stats["rounds"] == 0


def test_terminal_samples_stay_in_bounds_and_near_the_terminals(track_planner):
    ow, world, planner = track_planner
    s, g = np.array([1.9, 1.9, 0.05]), np.array([-1.95, -1.9, 1.95])      # next to the bounds: samples are clipped
    pts = planner._terminal_samples(s, g, 1000, 1)
    assert pts.shape == (124, 3)
    assert (pts >= planner.lo).all() and (pts <= planner.hi).all()
    d = np.minimum(np.abs(pts - s).max(axis=1), np.abs(pts - g).max(axis=1))
    assert (d <= planner.range + 1e-12).all()
    assert (np.abs(pts[0::4] - s).max(axis=1) <= planner.range).all() and (np.abs(pts[3::4] - g).max(axis=1) <= 0.25 * planner.range + 1e-12).all()
