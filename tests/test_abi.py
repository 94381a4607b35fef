"""The C-ABI library loads on a CPU-only box and exports exactly what include/gpushare_b200.h declares.
No compute call is made here."""
import ctypes
import os
import re
import subprocess

import pytest

import gsx
from gsx import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADERS = [os.path.join(ROOT, "include", "gpushare_b200.h"), os.path.join(ROOT, "include", "gpushare_ext.h")]


def declared():
    names = set()
    for h in HEADERS:
        src = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names |= set(re.findall(r"\b(gs_[a-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_header_symbols_are_exported():
    lib = capi.load()
    names = declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in gpushare_b200.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", capi.library_path()], capture_output=True, text=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l)
    assert exported == names, "exports and header disagree"
    assert sorted(capi.EXPORTS + gsx.ext.EXPORTS) == names


def test_struct_layouts_match_header(tmp_path):
    prog = tmp_path / "sz.c"
    prog.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "gpushare_b200.h"\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n",'
                    'sizeof(gs_node),sizeof(gs_pod_record),sizeof(gs_pod_event),sizeof(gs_bind_req),'
                    'offsetof(gs_node,node_class),offsetof(gs_pod_record,dev),offsetof(gs_pod_event,op),offsetof(gs_bind_req,counted));return 0;}')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), str(prog), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [capi.NODE_DT.itemsize, capi.POD_DT.itemsize, capi.EVENT_DT.itemsize, capi.BIND_DT.itemsize,
            capi.NODE_DT.fields["node_class"][1], capi.POD_DT.fields["dev"][1], capi.EVENT_DT.fields["op"][1],
            capi.BIND_DT.fields["counted"][1]]
    assert got == want


def test_abi_version_and_no_cpu_fallback():
    lib = capi.load()
    assert lib.gs_abi_version() == 2
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(gsx.GpushareError) as e:
            gsx.Snapshot(16, 4, 16)
        assert "no CPU fallback" in str(e.value)


def test_product_package_does_not_import_the_oracle():
    """The product path may not import, link or dlopen anything under oracle/."""
    pkg = os.path.join(ROOT, "gpushare-scheduler-extender_b200")
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|gpushare_oracle|orc_[a-z_]+\(|oracle/", re.M)
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".map")):
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert not pat.search(txt), f"{f} reaches into oracle/"
