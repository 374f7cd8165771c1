"""Import shim: ``import online_traj_planner`` (the reference's pybind module name, src/pybind.cpp:10) loads
efficient-path-planner_b200/online_traj_planner.py."""
import epp_b200  # noqa: F401  (registers the package under an importable name)
from epp_b200.online_traj_planner import *  # noqa: F401,F403
from epp_b200.online_traj_planner import BatchPlanner, MatrixXd, OnlineTrajGenerator, PathPlanner, Vector3d  # noqa: F401
