"""the PathWriter dumps (path_<k>.txt, checkpoints.txt, gates.txt, obstacles.txt) are byte-equal to
what the reference's own src/PathWriter.cpp writes — against the committed files of tests/golden/path_writer/
(tests/golden/make_path_writer_golden.py) and, where oracle/_ref is built, against the live reference code."""
import filecmp
import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
NAMES = ["checkpoints.txt", "gates.txt", "obstacles.txt", "path_0.txt", "path_1.txt"]


def test_dumps_equal_reference_golden(tmp_path):
    import online_traj_planner as otp
    from make_path_writer_golden import script
    out = tmp_path / "path_segments"
    script(otp.PathWriter(str(out)))
    assert sorted(os.listdir(out)) == NAMES
    for name in NAMES:
        assert (out / name).read_bytes() == open(os.path.join(HERE, "golden", "path_writer", name), "rb").read(), name


def test_existing_folder_is_emptied_of_files_only(tmp_path):
    import online_traj_planner as otp
    out = tmp_path / "dumps"
    (out / "keep_dir").mkdir(parents=True)
    (out / "old.txt").write_text("x")
    w = otp.PathWriter(str(out))
    assert sorted(os.listdir(out)) == ["keep_dir"]
    w.writePath([[0, 0, 0]])
    w.writePath([[1, 2, 3]])
    assert sorted(os.listdir(out)) == ["keep_dir", "path_0.txt", "path_1.txt"]
    assert (out / "path_1.txt").read_text() == "1 2 3 \n"
    otp.PathWriter(None).writePath([[0, 0, 0]])       # disabled writer: no folder, no files


def test_dumps_equal_live_reference(tmp_path):
    ref_py = pytest.importorskip("ref_py")
    if not ref_py.available():
        pytest.skip("oracle/_ref/libepp_ref.so not built (needs /root/reference)")
    import online_traj_planner as otp
    from make_path_writer_golden import script
    try:
        ref_writer = ref_py.RefPathWriter(str(tmp_path / "ref"))
    except AttributeError:
        pytest.skip("prebuilt oracle/_ref predates the PathWriter entry points")
    script(ref_writer)
    script(otp.PathWriter(str(tmp_path / "ours")))
    match, mismatch, errors = filecmp.cmpfiles(tmp_path / "ref", tmp_path / "ours", NAMES, shallow=False)
    assert sorted(match) == NAMES and not mismatch and not errors
