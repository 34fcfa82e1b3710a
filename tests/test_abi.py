"""CPU tier: the C-ABI library loads, exports every symbol include/vsp_b200.h declares, refuses to work without a
GPU (no CPU fallback), and its host-only entry point (domain selection) agrees with the oracle."""
import ctypes as C
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from vote_saver_lib_check import declared_symbols, exported_symbols  # noqa: E402

from oracle import cref


def test_library_exports_every_declared_symbol(vsp):
    decl = declared_symbols()
    assert len(decl) >= 36
    exp = set(exported_symbols(vsp.library_path()))
    assert [s for s in decl if s not in exp] == []
    # nothing but the vsp_ C ABI leaks out of the library
    assert [s for s in exp if not s.startswith("vsp_")] == []


def test_python_binding_types_every_declared_symbol(vsp):
    from importlib import import_module
    lib_mod = import_module("vote-saver-protocol_b200._lib")
    bound = {name for name, _r, _a in lib_mod.SYMBOLS}
    assert sorted(bound) == declared_symbols()
    lib_mod.load()
    assert lib_mod.load().vsp_version().startswith(b"vsp_b200")


def test_no_cpu_fallback_without_device(vsp):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the failure path is only observable on the CPU box")
    with pytest.raises(vsp.VspError) as e:
        vsp.Context(0)
    assert e.value.code == -2  # VSP_E_NO_DEVICE


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "vote-saver-protocol_b200")
    for dirpath, _d, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("the oracle", "").replace("against the oracle", ""), f


@pytest.mark.parametrize("m", [2, 3, 5, 6, 8, 24, 100, 1000, 1100, 5000, (1 << 20) - 31 + 31, (1 << 20) + 1, 1 << 32])
def test_domain_plan_matches_oracle(vsp, m):
    assert vsp.domain_plan(m) == cref.domain_plan(m)


def test_domain_plan_rejects_degenerate_sizes(vsp):
    for m in (0, 1):
        with pytest.raises(ValueError):
            vsp.domain_plan(m)


def test_header_is_plain_c_and_the_library_links_from_c(vsp, tmp_path):
    """include/vsp_b200.h under gcc -std=c11 -pedantic, and tests/cpp/c_abi_check.c linked against the .so and run: host-only
    entry points from C, VSP_E_NO_DEVICE (or a working context on a GPU box) from vsp_ctx_create"""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("no gcc")
    lib_dir = os.path.dirname(vsp.library_path())
    exe = str(tmp_path / "c_abi_check")
    subprocess.check_call([gcc, "-std=c11", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                           "-o", exe, os.path.join(ROOT, "tests", "cpp", "c_abi_check.c"), "-L", lib_dir, "-lvsp_b200",
                           "-Wl,-rpath," + lib_dir])
    run = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert run.returncode == 0 and run.stdout.startswith("OK"), run.stdout + run.stderr
