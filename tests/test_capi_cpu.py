"""No-GPU checks of the boundary: the C-ABI library builds, loads, and exports every symbol include/b3d.h declares; the ctypes
structs match the C layout; the product package never imports the oracle."""
import ctypes as C
import os
import re
import subprocess

import numpy as np

from conftest import ROOT, pkg


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "b3d.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b3d_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    capi = pkg("_capi")
    capi.build()
    lib = capi.lib()
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b3d.h but not exported"
    assert b"sm_100a" in lib.b3d_version()
    assert lib.b3d_last_error() is not None


def test_struct_layouts_match_header():
    capi = pkg("_capi")
    # compile a tiny C program against the header and compare sizeof / offsetof with the ctypes mirrors
    prog = r'''
#include <stdio.h>
#include <stddef.h>
#include "b3d.h"
int main(void) {
  printf("%zu %zu %zu %zu %zu %zu %zu\n", sizeof(b3d_model_desc), sizeof(b3d_fit_opts), sizeof(b3d_kp), sizeof(b3d_iter_record),
         offsetof(b3d_model_desc, v_template), offsetof(b3d_fit_opts, fx), offsetof(b3d_iter_record, flags));
  return 0; }
'''
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.c"), "w").write(prog)
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "t"), os.path.join(d, "t.c")], check=True)
        out = subprocess.run([os.path.join(d, "t")], check=True, capture_output=True, text=True).stdout.split()
    got = [int(x) for x in out]
    want = [C.sizeof(capi.ModelDesc), C.sizeof(capi.FitOpts), capi.KP_DTYPE.itemsize, capi.ITER_DTYPE.itemsize,
            capi.ModelDesc.v_template.offset, capi.FitOpts.fx.offset, capi.ITER_DTYPE.fields["flags"][1]]
    assert got == want


def test_index_tables_bit_exact(synth):
    api_src = open(os.path.join(ROOT, "3dbodyanimation_b200", "api.py")).read()
    # tables as in include/Utils.h:18-23 (USE_SMPL: 15 initialisers + two value-initialised zeros)
    mp = [-1, 23, 24, -1, 25, 26, -1, 27, 28, -1, 31, 32, -1, -1, -1, 0, 11, 12, 13, 14, 15, 16, -1, -1]
    use = [1, 2, 4, 5, 7, 8, 10, 11, 15, 16, 17, 18, 19, 20, 21, 0, 0]
    assert synth.MP_MAP.tolist() == mp and synth.USE_SMPL.tolist() == use
    assert str(mp) in api_src and str(use) in api_src


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "3dbodyanimation_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle_api" not in txt and "liboracle" not in txt and '#include "../../oracle' not in txt, f


def test_reference_signature_shim_builds_and_fails_loudly_without_gpu(tmp_path):
    """include/b3d_shim.hpp (the reference's C++ signatures on the C-ABI) compiles with the host compiler and links against the
    library; without a CUDA device the calls raise (B3D_ECUDA) instead of falling back to anything."""
    import shutil
    import subprocess
    import torch
    cxx = shutil.which("g++")
    if cxx is None:
        pytest.skip("no g++")
    libdir = os.path.join(ROOT, "3dbodyanimation_b200")
    exe = str(tmp_path / "shim_test")
    res = subprocess.run([cxx, "-std=c++14", "-Wall", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "shim_compile.cpp"), "-o", exe,
                          "-L" + libdir, "-l:lib3dba_b200.so", "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    run = subprocess.run([exe], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert run.returncode == 0, run.stdout + run.stderr
    else:
        assert run.returncode == 3 and "no CPU fallback" in run.stdout, run.stdout + run.stderr
