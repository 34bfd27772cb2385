"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/gmlrt.h declares; blob layout constants agree between header and host lowering."""
import os
import re

import pytest

from go_raytracer_b200 import cuda, lowering, scenes
from util import scene_blobs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "gmlrt.h")).read()
    declared = set(re.findall(r"\b(gmlrt_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(cuda.EXPORTS)
    lib = cuda.load()
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.gmlrt_blob_version() == lowering.VERSION
    assert b"sm_100a" in lib.gmlrt_build_info()


def test_struct_sizes_match_header():
    hdr = open(os.path.join(ROOT, "include", "gmlrt.h")).read()
    for name, size in (("gmlrt_blob_header", lowering.HEADER_BYTES), ("gmlrt_node", lowering.NODE_BYTES),
                       ("gmlrt_light", lowering.LIGHT_BYTES), ("gmlrt_surface", lowering.SURFACE_BYTES)):
        m = re.search(r"\}\s*%s;\s*/\*\s*(\d+) bytes" % name, hdr)
        assert m and int(m.group(1)) == size


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(cuda.GmlrtError) as ei:
        cuda.Context()
    assert ei.value.code == -4          # GMLRT_E_NO_DEVICE


def test_lowering_blob_roundtrip():
    (_, blob), = scene_blobs("cube")
    d = lowering.describe(blob)
    assert d["header"]["magic"] == lowering.MAGIC and d["header"]["total_bytes"] == len(blob)
    kinds = [n["kind"] for n in d["nodes"]]
    assert kinds.count(17) == 1 and kinds.count(20) == 2 and kinds.count(0) == 2
    code_surf = [s for s in d["surfaces"] if s["kind"] == 1]
    assert len(code_surf) == 1
    m = code_surf[0]["material"]
    assert m[5] == 0.0 and m[3] == 0.9 and m[0] != m[0]     # transparency known 0, refl = ks, colour per hit (NaN)


def test_synthetic_scene_generators_are_deterministic():
    a = scenes.to_blob(scenes.spheres_lattice(64, 36, grid=(4, 2, 2)))
    b = scenes.to_blob(scenes.spheres_lattice(64, 36, grid=(4, 2, 2)))
    assert a == b
    d = lowering.describe(a)
    assert sum(1 for n in d["nodes"] if n["kind"] == 16) == 16 and [l["kind"] for l in d["lights"]] == [0, 1]
    d = lowering.describe(scenes.to_blob(scenes.csg(64, 36, groups=(2, 2))))
    assert sum(1 for n in d["nodes"] if n["kind"] in (1, 2)) == 12 and all(l["kind"] == 2 for l in d["lights"])


def test_stats_mirrors_match_header_size():
    """gmlrt_stats is filled by C code through a pointer: both ctypes mirrors must be as large as the C struct."""
    import ctypes
    import oracle
    hdr = open(os.path.join(ROOT, "include", "gmlrt.h")).read()
    body = re.search(r"typedef struct gmlrt_stats \{(.*?)\} gmlrt_stats;", hdr, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    size = 0
    for decl in body.split(";"):
        m = re.match(r"\s*(double|uint64_t|uint32_t)\s+(.*)", decl.strip(), re.S)
        if not m:
            continue
        width = {"double": 8, "uint64_t": 8, "uint32_t": 4}[m.group(1)]
        for name in m.group(2).split(","):
            dim = re.search(r"\[(\d+)\]", name)
            size += width * (int(dim.group(1)) if dim else 1)
    assert ctypes.sizeof(cuda.Stats) == size == ctypes.sizeof(oracle.Stats)
