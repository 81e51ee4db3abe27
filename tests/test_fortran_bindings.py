"""CPU test: fortran/fempar_b200_bindings.f90 (the ISO_C_BINDING module a FEMPAR maintainer adds) is generated from
include/fempar_b200.h and declares every entry point of it; the hand-written type extensions only call declared entries."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bindings_module_is_up_to_date_and_complete():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "fortran", "gen_bindings.py"), "--check"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    hdr = open(os.path.join(ROOT, "include", "fempar_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(fb2_[a-z0-9_]+)\s*\(", hdr)))
    mod = open(os.path.join(ROOT, "fortran", "fempar_b200_bindings.f90")).read()
    bound = set(re.findall(r"name='(fb2_[a-z0-9_]+)'", mod))
    assert set(declared) == bound and len(bound) >= 70


def test_type_extensions_call_only_declared_entry_points():
    mod = open(os.path.join(ROOT, "fortran", "fempar_b200_bindings.f90")).read()
    bound = set(re.findall(r"name='(fb2_[a-z0-9_]+)'", mod))
    types = {"fb2_ils_params_t", "fb2_ils_ext_params_t"}
    used = set()
    for f in os.listdir(os.path.join(ROOT, "fortran")):
        if f.startswith("b200_") and f.endswith(".f90"):
            src = open(os.path.join(ROOT, "fortran", f)).read()
            src = "\n".join(ln.split("!")[0] for ln in src.splitlines())       # drop comments
            used |= set(re.findall(r"\b(fb2_[a-z0-9_]+)\b", src))
    assert used - types <= bound, sorted(used - types - bound)
    assert len(used) >= 30
