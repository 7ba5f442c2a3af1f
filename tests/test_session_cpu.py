"""Stored-session formats: the product's printf/strtod coding (csrc/vx_session.cu) against the oracle's
iostream coding (oracle/session.hpp), byte for byte, plus round trips and the reference's edge-reversal rule.  Host code only:
runs without a GPU."""
import ctypes as C
import os

import numpy as np
import pytest

import synth


@pytest.fixture(scope="module")
def pkg():
    import __graft_entry__ as ge
    ge.build_product()
    p = ge.load_package(); p.load_library()
    return p


@pytest.fixture(scope="module")
def olib():
    import __graft_entry__ as ge
    ge.build_oracle()
    import oracle_bind
    return oracle_bind.lib()


def states(pkg, n, seed, full=True):
    rng = np.random.Generator(np.random.PCG64(seed))
    R, p = synth.random_poses(seed, n, rot_sigma=1.5, tra_sigma=40.0)
    st = np.zeros(n, dtype=pkg.scan_state_dtype)
    st["t"] = 1.7e9 + np.cumsum(rng.uniform(0.05, 0.15, n))
    st["p"] = p; st["R"] = np.transpose(R, (0, 2, 1)).reshape(n, 9)
    st["v"] = rng.normal(0, 1, (n, 3)); st["bg"] = rng.normal(0, 1e-3, (n, 3)); st["ba"] = rng.normal(0, 1e-2, (n, 3))
    st["g"] = np.array([0, 0, -9.8]) + rng.normal(0, 0.05, (n, 3)); st["v6"] = rng.uniform(1e-6, 1e-3, (n, 6))
    return st


def test_quaternion_conversions_match_eigen_formulas(pkg, olib):
    lib = pkg.load_library()
    rng = np.random.Generator(np.random.PCG64(5))
    # all four branches of Quaterniond(Matrix3d): positive trace and each dominant diagonal entry
    Rs = list(synth.random_poses(3, 40, rot_sigma=2.5)[0])
    for ax in range(3):
        w = np.zeros(3); w[ax] = np.pi - 1e-3
        Rs.append(synth.rodrigues(w))
    for R in Rs:
        Rc = np.ascontiguousarray(R.T.reshape(9)); qa = np.zeros(4); qb = np.zeros(4)
        lib.vx_R_to_quat(Rc.ctypes.data_as(C.c_void_p), qa.ctypes.data_as(C.c_void_p))
        olib.orc_R_to_quat(Rc.ctypes.data_as(C.c_void_p), qb.ctypes.data_as(C.c_void_p))
        assert np.array_equal(qa, qb) and abs(np.linalg.norm(qa) - 1) < 1e-12
        Ra = np.zeros(9); Rb = np.zeros(9)
        lib.vx_quat_to_R(qa.ctypes.data_as(C.c_void_p), Ra.ctypes.data_as(C.c_void_p))
        olib.orc_quat_to_R(qa.ctypes.data_as(C.c_void_p), Rb.ctypes.data_as(C.c_void_p))
        assert np.array_equal(Ra, Rb) and np.abs(Ra.reshape(3, 3).T - R).max() < 1e-12


def test_lidarstate_same_bytes_and_round_trip(pkg, olib, tmp_path):
    st = states(pkg, 120, 11)
    a = str(tmp_path / "a.txt"); b = str(tmp_path / "b.txt")
    pkg.write_lidarstate(a, st)
    assert olib.orc_write_lidarstate(b.encode(), st.ctypes.data_as(C.c_void_p), C.c_int64(st.shape[0])) == 0
    assert open(a, "rb").read() == open(b, "rb").read()
    first = open(a).readline().split(" ")
    assert len(first) == 26 and len(first[0].split(".")[1]) == 6 and len(first[1].split(".")[1]) == 7
    got = pkg.read_lidarstate(a)
    ref = np.zeros(200, dtype=pkg.scan_state_dtype); n = C.c_int64(0)
    assert olib.orc_read_lidarstate(a.encode(), ref.ctypes.data_as(C.c_void_p), C.c_int64(200), C.byref(n)) == 0
    assert n.value == 120 and got.tobytes() == ref[:120].tobytes()
    assert np.abs(got["p"] - st["p"]).max() < 1e-7 and np.abs(got["t"] - st["t"]).max() < 1e-6 and np.abs(got["R"] - st["R"]).max() < 1e-6
    # 8-field (t p q) and 20-field lines keep the IMUST defaults for the rest (voxelslam.hpp:290-303)
    short = str(tmp_path / "short.txt")
    with open(short, "w") as f:
        f.write("1.5 1 2 3 0 0 0 1\n")
        f.write("2.5 1 2 3 0 0 0.7071068 0.7071068 0.1 0.2 0.3 0 0 0 0 0 0 0 0 -9.81\n")
    s2 = pkg.read_lidarstate(short)
    assert s2.shape[0] == 2 and np.array_equal(s2["R"][0], np.eye(3).reshape(9)) and np.array_equal(s2["g"][0], [0, 0, -9.8])
    assert np.array_equal(s2["v"][1], [0.1, 0.2, 0.3]) and s2["g"][1][2] == -9.81 and np.all(s2["v6"] == 0)
    with pytest.raises(pkg.VxError):
        pkg.read_lidarstate(str(tmp_path / "missing.txt"))
    with open(short, "w") as f:
        f.write("1.5 1  2 3 0 0 0 1\n")          # an empty field: std::stod throws in the reference
    with pytest.raises(pkg.VxError):
        pkg.read_lidarstate(short)


def test_pcd_same_bytes_and_round_trip(pkg, olib, tmp_path):
    rng = np.random.Generator(np.random.PCG64(3))
    for n in (0, 1, 5000):
        pts = rng.normal(0, 20, (n, 4)).astype(np.float32)
        a = str(tmp_path / ("a%d.pcd" % n)); b = str(tmp_path / ("b%d.pcd" % n))
        pkg.write_pcd_xyzi(a, pts)
        assert olib.orc_write_pcd_xyzi(b.encode(), pts.ctypes.data_as(C.c_void_p), C.c_int64(n)) == 0
        raw = open(a, "rb").read()
        assert raw == open(b, "rb").read()
        assert raw.startswith(b"# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z intensity\nSIZE 4 4 4 4\nTYPE F F F F\nCOUNT 1 1 1 1\n")
        assert raw.endswith(pts.tobytes()) and (b"\nPOINTS %d\nDATA binary\n" % n) in raw
        assert np.array_equal(pkg.read_pcd_xyzi(a).view(np.uint32), pts.view(np.uint32))
    # an ascii file with another field order and an extra field, no intensity -> intensity reads as 0
    c = str(tmp_path / "c.pcd")
    with open(c, "w") as f:
        f.write("# comment\nVERSION 0.7\nFIELDS ring z y x\nSIZE 2 4 4 4\nTYPE U F F F\nCOUNT 1 1 1 1\nWIDTH 2\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS 2\nDATA ascii\n")
        f.write("7 3.5 2.5 1.5\n9 -3 -2 -1\n")
    assert np.array_equal(pkg.read_pcd_xyzi(c), np.array([[1.5, 2.5, 3.5, 0], [-1, -2, -3, 0]], dtype=np.float32))
    d = str(tmp_path / "d.pcd")
    with open(d, "w") as f:
        f.write("VERSION 0.7\nFIELDS x y z\nSIZE 4 4 4\nTYPE F F F\nCOUNT 1 1 1\nWIDTH 1\nHEIGHT 1\nPOINTS 1\nDATA binary_compressed\n")
    with pytest.raises(pkg.VxError):
        pkg.read_pcd_xyzi(d)


def test_edges_same_bytes_reversal_and_absent_lines(pkg, olib, tmp_path):
    rng = np.random.Generator(np.random.PCG64(8))
    names = ["siteA", "siteB", "siteC"]
    n = 7
    R, p = synth.random_poses(21, n, rot_sigma=1.0, tra_sigma=15.0)
    e = np.zeros(n, dtype=pkg.pgo_edge_dtype)
    e["m1"] = [0, 0, 0, 1, 1, 0, 2]; e["m2"] = [0, 1, 2, 1, 2, 2, 2]
    e["id1"] = rng.integers(0, 500, n); e["id2"] = rng.integers(0, 500, n)
    e["rot"] = np.transpose(R, (0, 2, 1)).reshape(n, 9); e["tra"] = p; e["cov"] = 1e-6
    a = str(tmp_path / "edge_a.txt"); b = str(tmp_path / "edge_b.txt")
    pkg.write_edges(a, names, e)
    quat = np.zeros((n, 4))
    for i in range(n):
        olib.orc_R_to_quat(np.ascontiguousarray(e["rot"][i]).ctypes.data_as(C.c_void_p), quat[i].ctypes.data_as(C.c_void_p))
    n1 = (C.c_char_p * n)(*[names[k].encode() for k in e["m1"]]); n2 = (C.c_char_p * n)(*[names[k].encode() for k in e["m2"]])
    id1 = np.ascontiguousarray(e["id1"]); id2 = np.ascontiguousarray(e["id2"]); tra = np.ascontiguousarray(e["tra"])
    assert olib.orc_write_edges(b.encode(), n1, n2, id1.ctypes.data_as(C.c_void_p), id2.ctypes.data_as(C.c_void_p), tra.ctypes.data_as(C.c_void_p),
                                quat.ctypes.data_as(C.c_void_p), C.c_int64(n)) == 0
    assert open(a, "rb").read() == open(b, "rb").read()
    got, absent = pkg.read_edges(a, names, "siteC")
    assert absent == "" and got.shape[0] == n
    assert np.array_equal(got["m1"], e["m1"]) and np.array_equal(got["id2"], e["id2"]) and np.all(got["cov"] == 1e-6)
    assert np.abs(got["rot"] - e["rot"]).max() < 1e-6 and np.abs(got["tra"] - e["tra"]).max() < 1e-5
    # the loader knows the sessions in another order: a line whose names arrive reversed is inverted (voxelslam.cpp:244-250)
    got2, _ = pkg.read_edges(a, ["siteB", "siteA", "siteC"], "siteC")
    k = 1                                                    # line "siteA siteB": now mp = (1, 0) -> swapped
    assert (got2["m1"][k], got2["m2"][k], got2["id1"][k], got2["id2"][k]) == (0, 1, e["id2"][k], e["id1"][k])
    Rk = got["rot"][k].reshape(3, 3).T
    assert np.abs(got2["rot"][k].reshape(3, 3).T - Rk.T).max() < 1e-15 and np.abs(got2["tra"][k] + Rk.T @ got["tra"][k]).max() < 1e-12
    # only siteA loaded, current bag siteC: lines of A-A become edges, lines without siteC are kept verbatim, the rest dropped
    got3, absent3 = pkg.read_edges(a, ["siteA"], "siteC")
    lines = open(a).read().splitlines(True)
    assert got3.shape[0] == 1 and absent3 == "".join(l for l in lines if l.split()[:2] in (["siteA", "siteB"], ["siteB", "siteB"]))
    c = str(tmp_path / "edge_c.txt")
    pkg.write_edges(c, ["siteA"], got3, absent=absent3)
    assert open(c).read() == absent3 + lines[0]
    assert pkg.read_edges(str(tmp_path / "none.txt"), names, "x")[0].shape[0] == 0     # a missing edge.txt is an empty edge set
