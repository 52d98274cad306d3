"""CPU: the C-ABI library loads and exports every symbol include/rt_b200.h declares; without a GPU every compute
entry point fails loudly (there is no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import triangles_sdf_cpu_raytracing_b200 as rt
from triangles_sdf_cpu_raytracing_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = set()
    for h in sorted(os.listdir(os.path.join(ROOT, "include"))):
        text = open(os.path.join(ROOT, "include", h)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names |= set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", text))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    lib = rt.lib()
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rt_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes signature table out of sync with the header"


def test_struct_layouts_match_header():
    # rt_frame: 3+16+16 floats, 5 u32, i32, 3 floats, 3 i32, 1 float, 2 floats
    assert C.sizeof(rt.RtFrame) == 4 * (3 + 16 + 16 + 5 + 1 + 3 + 3 + 1 + 2 + 1)
    assert C.sizeof(rt.RtCounters) == 9 * 8


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(rt.RtError):
        rt.SDFGrid(np.zeros((4, 4, 4), np.float32))


def test_argument_validation_reports_errors():
    lib = rt.lib()
    h = C.c_void_p()
    assert lib.rt_scene_create_grid(None, 4, 4, 4, 0, C.byref(h)) == -1
    assert b"null" in lib.rt_last_error()
    assert lib.rt_scene_create_octree(None, 0, C.byref(h)) == -1
    assert lib.rt_set_math_mode(7) == -1
    assert lib.rt_set_sched_mode(-1) == -1
    assert lib.rt_scene_kind_of(None) == -1


def test_headers_are_plain_c(tmp_path):
    """The boundary is a C ABI: both headers compile as C99 with no C++ / CUDA / torch types, and a C program links
    against the library using only what they declare."""
    import shutil
    import subprocess
    cc = shutil.which("gcc") or shutil.which("cc")
    if not cc:
        pytest.skip("no C compiler")
    src = tmp_path / "abi.c"
    src.write_text("""
#include <stdio.h>
#include "rt_b200.h"
#include "rt_b200_host.h"
int main(void) {
  rt_frame f;
  rt_counters c;
  int n = -1;
  (void)f; (void)c;
  printf("%s %d %d %d\\n", rt_version(), (int)sizeof(rt_frame), (int)sizeof(rt_counters), rt_device_count(&n) == RT_OK ? n : -1);
  return rt_set_math_mode(RT_MATH_EXACT) == RT_OK && rt_set_sched_mode(RT_SCHED_COMPACT) == RT_OK &&
         rt_set_sched_mode(RT_SCHED_AUTO) == RT_OK ? 0 : 1;
}
""")
    exe = tmp_path / "abi"
    libdir = os.path.dirname(_lib.LIB_PATH)
    r = subprocess.run([cc, "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src),
                        "-o", str(exe), "-L", libdir, "-lrt_b200", f"-Wl,-rpath,{libdir}"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.split()[-3:-1] == [str(C.sizeof(rt.RtFrame)), str(C.sizeof(rt.RtCounters))]
