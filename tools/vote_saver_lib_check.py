"""Header <-> library export check: every function include/vsp_b200.h declares must be a dynamic symbol of the .so."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols(header=None):
    header = header or os.path.join(ROOT, "include", "vsp_b200.h")
    return sorted(set(re.findall(r"VSP_API[^;(]*?\b(vsp_[a-z0-9_]+)\s*\(", open(header).read())))


def exported_symbols(so_path):
    out = subprocess.check_output(["nm", "-D", "--defined-only", so_path], text=True)
    return sorted({ln.split()[-1] for ln in out.splitlines() if " T " in ln})


def check_exports(so_path):
    missing = [s for s in declared_symbols() if s not in set(exported_symbols(so_path))]
    if missing:
        raise RuntimeError("%s does not export: %s" % (so_path, ", ".join(missing)))
    return True
