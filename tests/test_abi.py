"""CPU suite: the C-ABI library loads, exports every symbol include/rte.h declares, agrees with the ctypes
mirror on struct layout, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from raytracingengine_b200 import _abi
from raytracingengine_b200.engine import LIB_PATH, SYMBOLS, RteError, load_library

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rte.h")


def header_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rte_[a-z0-9_]+)\s*\(", src)))


def test_library_is_built_in_tree():
    assert os.path.exists(LIB_PATH), "build raytracingengine_b200/librte_b200.so first (__graft_entry__.build())"
    assert os.path.dirname(LIB_PATH).endswith("raytracingengine_b200")


def test_every_declared_symbol_is_exported():
    lib = load_library()
    declared = header_functions()
    assert declared, "no functions parsed from include/rte.h"
    assert sorted(SYMBOLS) == declared, "engine.SYMBOLS and include/rte.h disagree"
    for name in declared:
        assert hasattr(lib, name), f"librte_b200.so does not export {name}"
    dyn = subprocess.run(["nm", "-D", "--defined-only", LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (rte_[a-z0-9_]+)", dyn))
    assert exported == set(declared), f"extra/missing exports: {exported ^ set(declared)}"
    assert lib.rte_version() == _abi.RTE_VERSION


def test_struct_layouts_match_the_header(tmp_path):
    """sizeof/offsetof from a C compile of include/rte.h == the ctypes / numpy mirrors in _abi.py."""
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "rte.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                    "sizeof(rte_material),sizeof(rte_sphere),sizeof(rte_plane),sizeof(rte_triangle),sizeof(rte_model),sizeof(rte_light),"
                    "sizeof(rte_camera),sizeof(rte_scene_desc),sizeof(rte_render_opts),sizeof(rte_frame),sizeof(rte_stats),sizeof(rte_hit),"
                    "offsetof(rte_render_opts,stripe_rows),offsetof(rte_hit,prim_id));return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)], check=True)
    got = list(map(int, subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()))
    want = [_abi.material_dt.itemsize, _abi.sphere_dt.itemsize, _abi.plane_dt.itemsize, _abi.triangle_dt.itemsize, _abi.model_dt.itemsize,
            _abi.light_dt.itemsize, C.sizeof(_abi.Camera), C.sizeof(_abi.SceneDesc), C.sizeof(_abi.RenderOpts), C.sizeof(_abi.Frame),
            C.sizeof(_abi.Stats), _abi.hit_dt.itemsize, _abi.RenderOpts.stripe_rows.offset, _abi.hit_dt.fields["prim_id"][1]]
    assert got == want
    # Material keeps the reference's 56-byte layout (SURVEY §8(c) struct-size KAT)
    assert _abi.material_dt.itemsize == 56 and _abi.light_dt.itemsize == 56


def test_header_is_plain_c():
    r = subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-fsyntax-only", "-x", "c", HEADER], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_no_cpu_fallback_without_a_device():
    """Without a GPU every compute entry point must fail loudly; with one, this test only checks create/destroy."""
    lib = load_library()
    ctx = C.c_void_p()
    rc = lib.rte_create(0, C.byref(ctx))
    if lib.rte_device_count() == 0:
        assert rc == _abi.RTE_ERR_NO_DEVICE and not ctx.value
        assert b"no CPU path" in lib.rte_last_error(None)
        from raytracingengine_b200.engine import Engine
        with pytest.raises(RteError):
            Engine(0)
        # NULL-ctx calls are argument errors, never silent successes
        assert lib.rte_upload_scene(None, None) == _abi.RTE_ERR_INVALID
        assert lib.rte_render(None, None, None) == _abi.RTE_ERR_INVALID
        assert lib.rte_tonemap(None, None, 0, 127, 0.0, 0.0, None) == _abi.RTE_ERR_INVALID
    else:
        assert rc == 0
        lib.rte_destroy(ctx)


def test_write_ppm_needs_no_device_and_matches_the_reference_writer(tmp_path, port_oracle):
    from raytracingengine_b200.engine import write_ppm
    rng = np.random.default_rng(3)
    rgb = rng.integers(0, 256, size=(7 * 5, 3), dtype=np.uint8)
    a, b = tmp_path / "a.ppm", tmp_path / "b.ppm"
    write_ppm(str(a), rgb, 7, 5)
    port_oracle.write_ppm(str(b), rgb, 7, 5)
    assert a.read_bytes() == b.read_bytes() == b"P6\n7 5\n255\n" + rgb.tobytes()
    with pytest.raises(RteError) as e:
        write_ppm(str(tmp_path / "no_such_dir" / "x.ppm"), rgb, 7, 5)
    assert e.value.status == _abi.RTE_ERR_IO and "Could not open file for writing" in str(e.value)  # RE/Image.cpp:15
    with pytest.raises(ValueError):
        write_ppm(str(a), rgb, 8, 5)


def test_product_package_never_touches_the_oracle():
    """A product path that routes through oracle/ would void every parity claim."""
    pkg = os.path.join(ROOT, "raytracingengine_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "pyoracle" not in src and "oracle_port" not in src and "liboracle" not in src and "libref_oracle" not in src, f
    libs = subprocess.run(["ldd", LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in libs


def test_write_png_decodes_to_the_same_pixels(tmp_path):
    """rte_write_png (row N4: what main() gets from ffmpeg): a valid PNG — signature, IHDR, one IDAT whose zlib stream
    inflates to filter-0 scanlines of the input pixels, IEND, all CRCs right.  Large enough for several stored blocks."""
    import struct
    import zlib
    from raytracingengine_b200.engine import RteError, write_png
    W, H = 301, 97  # 88 KB of scanlines: two 65535-byte stored blocks
    rgb = np.random.default_rng(3).integers(0, 256, (H, W, 3), dtype=np.uint8)
    path = tmp_path / "a.png"
    write_png(str(path), rgb, W, H)
    raw = path.read_bytes()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(raw):
        (n,), typ = struct.unpack(">I", raw[pos:pos + 4]), raw[pos + 4:pos + 8]
        body = raw[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])
        assert zlib.crc32(typ + body) == crc, typ
        chunks.append((typ, body))
        pos += 12 + n
    assert [c[0] for c in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    assert struct.unpack(">IIBBBBB", chunks[0][1]) == (W, H, 8, 2, 0, 0, 0)
    lines = np.frombuffer(zlib.decompress(chunks[1][1]), dtype=np.uint8).reshape(H, 1 + 3 * W)
    assert not lines[:, 0].any()
    assert np.array_equal(lines[:, 1:].reshape(H, W, 3), rgb)
    with pytest.raises(RteError):
        write_png(str(tmp_path / "no_such_dir" / "x.png"), rgb, W, H)
