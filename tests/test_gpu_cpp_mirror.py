"""GPU tests of the C++20 host mirror include/rte/engine.hpp (the drop-in boundary of SURVEY.md §8(b)).

* examples/reference_main = the body of the REFERENCE's own main() (RE/RaytracingEngine.cpp:216-330), compiled by
  __graft_entry__.build() against rte/engine.hpp with only the includes replaced (the source is read from /root/reference at build
  time and never committed; antiAliasingAmount defaults to 1 through RTE_ENGINE_DEFAULT_AA).  It loads box.obj FROM DISK, renders,
  runs tonemapAll and writes the seven PPMs: they must be the reference build's own within +-1 LSB (pow() in Blinn-Phong).
* examples/mirror_api exercises Scene::IntersectClosest / IntersectAnyBefore / CalculatePixelDepth, the operator functions, toColor,
  tonemap / tonemapAll on the known-answer vectors of tests/golden/kat.json."""
import json
import os
import subprocess

import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


def _exe(name):
    import __graft_entry__ as ge
    path = os.path.join(ROOT, "examples", name)
    if not os.path.exists(path):
        if name == "reference_main":
            if not ge.build_reference_main():
                pytest.skip("examples/reference_main is built where /root/reference exists (it travels to the GPU box prebuilt)")
        else:
            ge.build_cpp(name)
    return path


def test_reference_main_through_the_mirror(oracle, tmp_path):
    exe = _exe("reference_main")
    v, faces = scenes.box_mesh()
    scenes.write_obj(str(tmp_path / "box.obj"), v, faces)  # the stand-in of SURVEY D5, loaded from disk by LoadObject("box.obj")
    r = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "Temps de g" in r.stdout and "Image written to aces.ppm" in r.stdout  # upstream's own messages
    arrays, cam = scenes.config1()
    o = oracle.render(arrays, cam)
    want = oracle.tonemap_all(o["hdr"])
    differing = 0
    for op, name in enumerate(_abi.TM_NAMES):
        raw = (tmp_path / f"{name}.ppm").read_bytes()
        assert raw.startswith(b"P6\n1000 1000\n255\n") and len(raw) == 3000017  # RE/Image.cpp:18, the size of RE/output.ppm
        got = np.frombuffer(raw[17:], dtype=np.uint8).reshape(-1, 3)
        d = np.abs(got.astype(np.int16) - want[op].astype(np.int16))
        assert d.max() <= 1, f"{name}: {d.max()} LSB"
        differing += int(np.count_nonzero(d))
    assert differing <= 50, f"{differing} bytes differ (exact FP64 operators: only pow() in Blinn-Phong may move a byte)"


def test_default_scene_example(oracle, tmp_path):
    exe = _exe("default_scene")
    r = subprocess.run([exe, str(tmp_path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    arrays, cam = scenes.config1()
    want = oracle.tonemap_all(oracle.render(arrays, cam)["hdr"])
    for op, name in enumerate(_abi.TM_NAMES):
        got = np.frombuffer((tmp_path / f"{name}.ppm").read_bytes()[17:], dtype=np.uint8).reshape(-1, 3)
        assert np.abs(got.astype(np.int16) - want[op].astype(np.int16)).max() <= 1, name


def test_mirror_queries_and_operators_on_the_kats():
    with open(os.path.join(GOLD, "kat.json")) as f:
        kat = json.load(f)
    out = subprocess.run([_exe("mirror_api")], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    lines = {}
    for l in out.stdout.splitlines():
        k, _, rest = l.partition(" ")
        lines.setdefault(k, []).append(rest)

    def hit(key):
        t = lines[key][0].split()
        return None if t[0] == "miss" else (int(t[0]), int(t[1]), [float(x) for x in t[2:]])

    assert hit("sphere_front")[:2] == (1, 0) and hit("sphere_front")[2][0] == 4.0 and hit("sphere_front")[2][1:4] == [0.0, 0.0, -1.0]
    assert hit("sphere_inside")[2][0] == 1.0 and hit("sphere_inside")[2][1:4] == [0.0, 0.0, 1.0]   # far root, outward normal
    assert hit("sphere_away") is None and hit("sphere_tangent")[2][0] == 5.0
    assert lines["any_before"][0] == "1 0"
    assert hit("pixel_centre")[2][0] == 24.0 and hit("pixel_centre")[2][4:7] == [0.0, 0.0, -1.0]
    assert float(lines["normalized"][0]) == (24.0 - 0.0) / (200.0 - 0.0)
    assert lines["copy_equal"][0] == "1"
    assert hit("plane_hit")[:2] == (2, 0) and hit("plane_hit")[2][0] == 21.213203435596427  # dir = Vec3(0,-1,1).normalize() upstream
    assert hit("plane_on")[2][0] == 0.0 and hit("plane_parallel") is None
    assert hit("tri_front")[:2] == (3, 0) and hit("tri_front")[2][:4] == [3.0, 0.0, 0.0, 1.0]
    assert hit("tri_back")[2][0] == 3.0 and hit("tri_vertex")[2][0] == 3.0
    want = np.array(kat["tonemap"]["rgb8"])  # [op][vector][channel]
    for i in range(7):
        ops = np.array(lines["ops"][i].split()[1:], dtype=int).reshape(7, 3)
        assert np.array_equal(ops, want[:, i, :]), f"operator functions, vector {i}"
        a, _, one = lines["all"][i].partition("|")
        assert np.array_equal(np.array(a.split()[1:], dtype=int).reshape(7, 3), want[:, i, :]), f"tonemapAll, vector {i}"
        assert np.array_equal(np.array(one.split(), dtype=int), want[6, i, :]), f"tonemap, vector {i}"
    assert float(lines["luminance"][0]) == 0.5 * 0.2126 + 1 * 0.7152 + 2 * 0.0722
