"""CPU tests of the drop-in boundary: the shared library loads, exports every symbol the header declares,
and fails loudly (never falls back to a CPU path) when no CUDA device is present."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, load_system
from maniac_mc_b200 import lib


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "maniac_gpu.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(mmc_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    lib.build()
    L = C.CDLL(lib.SO_PATH)
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/maniac_gpu.h but not exported"
    assert sorted(lib.ABI_SYMBOLS) == syms


def test_struct_layouts_match_header():
    """ctypes mirrors must have the C struct sizes (x86-64 SysV): checked against a compiled probe."""
    import subprocess, tempfile
    src = '#include <stdio.h>\n#include "maniac_gpu.h"\nint main(){printf("%zu %zu %zu\\n", sizeof(mmc_desc), sizeof(mmc_move_record), sizeof(mmc_counters));return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "p.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "p"), os.path.join(d, "p.c")])
        a, b, c = map(int, subprocess.check_output([os.path.join(d, "p")]).split())
    from maniac_mc_b200.system import Counters, MmcDesc, MoveRecord
    assert (C.sizeof(MmcDesc), C.sizeof(MoveRecord), C.sizeof(Counters)) == (a, b, c)


def test_ewald_setup_host_entry_point():
    """mmc_ewald_setup is host arithmetic (prepare_utils.f90:169-214) and must agree with the oracle."""
    from oracle import oracle as orc
    for name in ("lj_gas", "h2o_gas", "zif8_h2o"):
        s = load_system(name)
        assert lib.ewald_setup(s.cutoff, s.ewald_tolerance, s.box_len) == orc.ewald_setup(s.cutoff, s.ewald_tolerance, s.box_len)


def test_no_cpu_fallback():
    """Without a GPU mmc_init must fail with a message; with one it must succeed."""
    import torch
    s = load_system("lj_gas")
    if torch.cuda.is_available():
        g = lib.ManiacGPU(s); g.close()
    else:
        with pytest.raises(lib.ManiacGpuError, match="no CUDA device|CUDA"):
            lib.ManiacGPU(s)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "maniac_mc_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("oracle/maniac_oracle.c:orc_philox_uniform", ""), f"{f} mentions the oracle"


def test_fortran_interface_matches_header():
    """fortran/maniac_gpu_iface.f90 cannot be compiled here (no Fortran compiler): check by text that its bind(C)
    mmc_desc lists the header's fields in the header's order with matching kinds, and that every bind(C, name=...)
    it declares is a symbol the library exports."""
    hdr = open(os.path.join(ROOT, "include", "maniac_gpu.h")).read()
    body = re.search(r"typedef struct mmc_desc \{(.*?)\} mmc_desc;", hdr, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    c_fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        m = re.match(r"(const\s+)?(int|double)\s*(\*?)\s*(\w+)(\[\d+\])?$", decl)
        assert m, decl
        kind = "ptr" if m.group(3) else m.group(2)
        c_fields.append((m.group(4), kind))
    f90 = open(os.path.join(ROOT, "fortran", "maniac_gpu_iface.f90")).read()
    tbody = re.search(r"type, bind\(C\) :: mmc_desc(.*?)end type mmc_desc", f90, re.S).group(1)
    f_fields = []
    for line in tbody.splitlines():
        line = line.split("!")[0].strip()
        if "::" not in line:
            continue
        typ, names = line.split("::")
        kind = "ptr" if "c_ptr" in typ else "int" if "c_int" in typ else "double"
        for n in re.findall(r"(\w+)(?:\(\d+\))?", names):
            f_fields.append((n, kind))
    assert f_fields == c_fields
    L = lib.load()
    names = re.findall(r'bind\(C, name="(\w+)"\)', f90)
    assert len(names) >= 15
    for n in names:
        assert hasattr(L, n), f"{n} bound in the Fortran interface but not exported"


def test_fortran_glue_uses_declared_names_only():
    """fortran/maniac_gpu_glue.f90 (compile-untested): every mmc_* it calls is declared in the interface module and
    exported, every mmc_desc component it sets is a header field, and it sets all of them."""
    glue = open(os.path.join(ROOT, "fortran", "maniac_gpu_glue.f90")).read()
    code = "\n".join(l.split("!")[0] for l in glue.splitlines())
    f90 = open(os.path.join(ROOT, "fortran", "maniac_gpu_iface.f90")).read()
    declared = set(re.findall(r'bind\(C, name="(\w+)"\)', f90)) | {"mmc_check", "mmc_desc"}
    used = set(re.findall(r"\b(mmc_\w+)", code))
    assert used and used <= declared, used - declared
    L = lib.load()
    for n in used - {"mmc_check", "mmc_desc"}:
        assert hasattr(L, n)
    hdr = open(os.path.join(ROOT, "include", "maniac_gpu.h")).read()
    body = re.sub(r"/\*.*?\*/", "", re.search(r"typedef struct mmc_desc \{(.*?)\} mmc_desc;", hdr, re.S).group(1), flags=re.S)
    fields = set(re.findall(r"(\w+)(?:\[\d+\])?\s*;", body))
    set_in_glue = set(re.findall(r"\bd%(\w+)", code))
    assert set_in_glue == fields, (set_in_glue ^ fields)
