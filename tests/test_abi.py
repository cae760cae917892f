"""The C-ABI library loads and exports every symbol include/b200rt.h declares; without a GPU every compute entry
point fails loudly (no CPU fallback).  No compute calls are made here."""
import ctypes as C
import os
import re

import pytest

import rust_raytracer_b200 as rt
from rust_raytracer_b200 import _ffi as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200rt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200rt_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound(built):
    syms = declared_symbols()
    assert len(syms) >= 40
    lib = C.CDLL(F.LIB_PATH)
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    assert sorted(F.PROTOTYPES) == syms  # the ctypes table covers exactly the header
    assert F.lib().b200rt_abi_version() == 1


def test_struct_sizes_match_the_header(built):
    assert C.sizeof(F.Prim) == 64 and C.sizeof(F.Material) == 16 and C.sizeof(F.Texture) == 48
    assert C.sizeof(F.Image) == 16 and C.sizeof(F.Instance) == 96 and C.sizeof(F.Hit) == 32
    assert C.sizeof(F.Perlin) == 256 * 16 + 3 * 256 * 4
    assert C.sizeof(F.Camera) == 24 * 4 and C.sizeof(F.Params) == 64 and C.sizeof(F.Stats) == 80


def test_no_cpu_fallback(built):
    n = C.c_int(-1)
    rc = F.lib().b200rt_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(rt.B200RTError) as e:
        rt.Context(0)
    assert e.value.status == F.ERR_CUDA and "no CPU fallback" in str(e.value)


def test_package_does_not_touch_the_oracle():
    """The product path may not import, link or execute anything under oracle/ (only tests/, smoke() and bench.py's CPU legs may)."""
    pkg = os.path.join(ROOT, "rust_raytracer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")) and f != "build.py":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                code = text.replace("the oracle/", "").replace("(oracle/oracle_scenes.inc)", "")  # comments may NAME the checker
                assert "liboracle" not in code and "oracle_ffi" not in code and "oracle_py" not in code and "oracle/" not in code, os.path.join(dirpath, f)


def test_write_png_round_trip(built, tmp_path):
    """RgbImage::save (main.rs:54): the dependency-free PNG writer decodes to the same pixels (PIL as the checker)."""
    import numpy as np
    from PIL import Image
    rng = np.random.default_rng(7)
    for (w, h) in ((1, 1), (37, 19), (300, 221)):  # 300x221x3+221 bytes: more than one 65535-byte stored block
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        path = str(tmp_path / ("t_%dx%d.png" % (w, h)))
        assert F.lib().b200rt_write_png(path.encode(), img.ctypes.data_as(C.c_void_p), w, h) == 0
        back = np.asarray(Image.open(path).convert("RGB"))
        assert back.shape == img.shape and (back == img).all()
    assert F.lib().b200rt_write_png(b"/nonexistent_dir/x.png", img.ctypes.data_as(C.c_void_p), 300, 221) == F.ERR_IO
    assert F.lib().b200rt_write_png(b"x.png", None, 4, 4) == F.ERR_INVALID


def test_cli_fails_loudly_without_a_device(built, tmp_path):
    """b200rt_render (the reference's main over the C ABI) has no CPU fallback either."""
    import subprocess
    n = C.c_int(-1)
    if F.lib().b200rt_device_count(C.byref(n)) == 0 and n.value > 0:
        pytest.skip("a CUDA device is present")
    from rust_raytracer_b200.build import CLI
    scene = os.path.join(ROOT, "data", "scene_10.json")
    r = subprocess.run([CLI, "--scene", scene, "--spp", "1", "--out", str(tmp_path / "o.png")], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr and not (tmp_path / "o.png").exists()
    r = subprocess.run([CLI, "--scene", "no_such_scene"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1 and "scene" in r.stderr


def test_rust_sys_crate_declares_the_same_symbols():
    """rust/b200rt-sys (shipped as source: no Rust toolchain here) must declare exactly the entry points of include/b200rt.h,
    and its repr(C) structs must have the sizes of the ctypes mirrors (same field lists)."""
    import re
    src = open(os.path.join(ROOT, "rust", "b200rt-sys", "src", "lib.rs")).read()
    declared = set(re.findall(r"pub fn (b200rt_\w+)\(", src))
    assert declared == set(F.PROTOTYPES.keys()), declared ^ set(F.PROTOTYPES.keys())
    size_of = {"u8": 1, "i32": 4, "u32": 4, "f32": 4, "u64": 8, "i64": 8, "f64": 8}
    for rust_name, ct in (("b200rt_camera", F.Camera), ("b200rt_params", F.Params), ("b200rt_stats", F.Stats), ("b200rt_hit", F.Hit),
                          ("b200rt_multi_stats", F.MultiStats), ("b200rt_sppm_params", F.SppmParams), ("b200rt_sppm_stats", F.SppmStats)):
        text = re.search(r"pub struct %s \{(.*?)\}" % rust_name, src, re.S).group(1)
        fields = re.findall(r"pub (\w+): (\[?\w+(?:; \d+\])?)", text)
        assert [f for f, _ in fields] == [f[0] for f in ct._fields_], rust_name
        total = 0
        for _, ty in fields:
            m = re.match(r"\[(\w+); (\d+)\]", ty)
            total += size_of[m.group(1)] * int(m.group(2)) if m else size_of[ty]
        assert total == C.sizeof(ct), (rust_name, total, C.sizeof(ct))  # no padding in any of them: naturally aligned field order
