"""Import shim: ``import polynomial_trajectory`` (the name the reference README installs the pybind module under) loads
efficient-path-planner_b200/online_traj_planner.py."""
import epp_b200  # noqa: F401  (registers the package under an importable name)
from epp_b200.online_traj_planner import *  # noqa: F401,F403
from epp_b200.online_traj_planner import BatchPlanner, MatrixXd, OnlineTrajGenerator, PathPlanner, Vector3d  # noqa: F401
