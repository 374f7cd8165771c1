"""Writes tests/golden/path_writer/*.txt with THE REFERENCE'S OWN PathWriter (src/PathWriter.cpp compiled unmodified
into oracle/_ref/libepp_ref.so; only Eigen's default stream format is a restatement, oracle/shim/Eigen/Dense).
Run where /root/reference exists: `make -C oracle ref && python tests/golden/make_path_writer_golden.py`."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))


def script(writer):
    """The same sequence of calls for the reference writer and for ours."""
    rng = np.random.default_rng(11)
    gates = np.array([[0.45, -1.0, 0, 0, 0, 2.35, 1], [1.0, -1.55, 0, 0, 0, -0.78, 0], [0.0, 0.5, 0, 0, 0, 0, 1],
                      [-0.5, -0.5, 0, 0, 0, 3.14, 0]])
    obstacles = np.array([[1.0, -0.5, 0, 0, 0, 0], [0.5, -1.5, 0, 0, 0, 0], [-0.5, 0.0, 0, 0, 0, 0], [0, 1.0, 0, 0, 0, 0]])
    for i, g in enumerate(gates):
        writer.updateGatePos(i, g)
    for i, o in enumerate(obstacles):
        writer.updateObstaclePos(i, o)
    pts = np.vstack([[[1.0, 1.0, 0.05], [0.0, -2.0, 0.5], [1e-7, -123456.789, 1e6], [1 / 3, 2e10, -0.0]],
                     rng.normal(size=(12, 3)) * 10.0 ** rng.integers(-6, 7, size=(12, 1))])
    writer.writeCheckpoints(pts[:7])
    writer.writePath(pts)
    writer.writePath(pts[::-1] * 0.37)
    writer.updateGatePos(2, gates[2] + [0.12345678, -0.09, 0, 0, 0, 1e-3, 0])
    writer.updateGatePos(1, rng.normal(size=7) * 1e-5)
    writer.writeCheckpoints(pts[3:])


if __name__ == "__main__":
    import ref_py
    out = os.path.join(HERE, "path_writer")
    script(ref_py.RefPathWriter(out))
    print(sorted(os.listdir(out)))
