"""The C-ABI library loads without a GPU and exports every symbol include/clrt.h declares; no compute calls here."""
import os
import re
import ctypes as C
import numpy as np
import pytest
from conftest import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "clrt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(clrt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from openclraytracer_b200 import _lib
    lib = _lib.load()
    names = declared_symbols()
    assert len(names) >= 40
    for name in names:
        assert hasattr(lib, name), f"{name} is declared in include/clrt.h but not exported"
        assert name in _lib.SIGNATURES, f"{name} has no ctypes signature in _lib.py"
    for name in _lib.SIGNATURES:
        assert name in names, f"{name} is bound in _lib.py but not declared in include/clrt.h"
    assert lib.clrt_abi_version() == 1


def test_no_cpu_fallback_without_gpu():
    """Without a CUDA device every entry point that needs one fails with an error message (never computes on the CPU)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from openclraytracer_b200 import api
    with pytest.raises(api.ClrtError, match="no CUDA device|no CPU fallback"):
        api.Context(0)


def test_product_package_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "openclraytracer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle_py" not in src and "libclrt_oracle" not in src and "libclrt_ref" not in src, f


def test_layout_sizes_match_header():
    """include/clrt_types.h compiled with gcc gives the sizes layouts.py asserts."""
    import subprocess, tempfile
    names = {"clrt_bvh_node": 48, "clrt_aabb": 32, "clrt_contact": 32, "clrt_ray": 48, "clrt_camera": 64, "clrt_matrix4": 48,
             "clrt_light": 16, "clrt_material": 96, "clrt_scene_header": 96, "clrt_model_header": 64, "clrt_mesh_header": 48,
             "clrt_grid_data": 64, "clrt_top_cell": 16, "clrt_triangle": 48, "clrt_uint2": 8}
    body = "".join(f'_Static_assert(sizeof({k}) == {v}, "{k}");\n' for k, v in names.items())
    with tempfile.TemporaryDirectory() as d:
        src = os.path.join(d, "t.c")
        open(src, "w").write(f'#include "clrt.h"\n{body}int main(void){{return 0;}}\n')
        subprocess.run(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), "-c", src, "-o", os.path.join(d, "t.o")], check=True)


def test_every_exported_symbol_is_documented():
    """INTEGRATION.md / DESIGN.md name every entry point of include/clrt.h (which reference interface it replaces, or that it is new)."""
    import re
    text = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "clrt.h")).read(), flags=re.S)
    names = sorted(set(re.findall(r"\b(clrt_[a-z0-9_]+)\s*\(", text)))
    docs = open(os.path.join(ROOT, "INTEGRATION.md")).read() + open(os.path.join(ROOT, "DESIGN.md")).read()
    assert len(names) > 60
    assert [n for n in names if n not in docs] == []


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="needs the reference tree")
def test_reference_citations_point_into_existing_lines():
    """Every `File.ext:first[-last]` citation in the headers, sources and documents names a file of the reference and lines it has."""
    import collections
    import re
    index = collections.defaultdict(list)
    for d, _, fs in os.walk("/root/reference"):
        for f in fs:
            index[f.lower()].append(os.path.join(d, f))
    own = {"clrt.h", "scene.cpp", "clrt_types.h", "clrt_host.h"}            # same-named files of this repository
    docs = ["DESIGN.md", "INTEGRATION.md", "include/clrt.h", "include/clrt_types.h", "include/clrt_host.h", "include/clrt/CLRayTracer.hpp",
            "host/Scene.cpp", "host/OpenCLRayTracerDemo.cpp", "oracle/clrt_oracle.cpp", "oracle/ref_bridge.cpp",
            "openclraytracer_b200/api.py", "openclraytracer_b200/frame.py"]
    csrc = os.path.join(ROOT, "openclraytracer_b200", "csrc")
    docs += [os.path.join("openclraytracer_b200", "csrc", f) for f in sorted(os.listdir(csrc)) if f.endswith((".cu", ".cuh"))]
    checked, bad = 0, []
    for doc in docs:
        text = open(os.path.join(ROOT, doc), errors="replace").read()
        for m in re.finditer(r"([A-Za-z_][A-Za-z0-9_]*\.(?:h|cpp|cl|txt|bat)):(\d+)(?:-(\d+))?", text):
            name, a, b = m.group(1), int(m.group(2)), int(m.group(3) or m.group(2))
            paths = index.get(name.lower())
            if not paths:
                if name.lower() not in own:
                    bad.append((doc, m.group(0), "no such file in the reference"))
                continue
            checked += 1
            lines = max(sum(1 for _ in open(p, errors="replace")) for p in paths)
            if a > b or b > lines:
                bad.append((doc, m.group(0), "the file has %d lines" % lines))
    assert checked > 250 and bad == []


def test_public_headers_compile_as_c99_and_cpp11(tmp_path):
    """The boundary is a C ABI: include/clrt.h and clrt_host.h must be plain C (what a cgo / JNI / ctypes binding would include);
    the C++ host mirror must be a clean C++11 header (the reference builds with VS2012)."""
    import subprocess
    c = tmp_path / "c.c"
    c.write_text("#include <clrt.h>\n#include <clrt_host.h>\nint main(void) { return clrt_abi_version() > 0 ? 0 : 0; }\n")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", str(c),
                    "-o", str(tmp_path / "c.o")], check=True)
    cpp = tmp_path / "h.cpp"
    cpp.write_text("#include <clrt/CLRayTracer.hpp>\nint main() { return 0; }\n")
    subprocess.run(["g++", "-std=c++11", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), "-c", str(cpp),
                    "-o", str(tmp_path / "h.o")], check=True)
