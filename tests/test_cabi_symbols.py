"""CPU: the C-ABI library loads without a GPU and exports every symbol include/mrgpu.h declares; calls that
need a device fail loudly (there is no CPU path)."""
import ctypes
import os
import re
import subprocess

import gpu_mapreduce_b200 as g

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "mrgpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mrgpu_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported():
    names = _declared()
    assert len(names) >= 30
    out = subprocess.check_output(["nm", "-D", "--defined-only", g.LIB_PATH], text=True)
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    missing = [n for n in names if n not in exported]
    assert not missing, missing
    lib = ctypes.CDLL(g.LIB_PATH)
    for n in names:
        assert hasattr(lib, n)


def test_every_declaration_cites_the_reference_or_says_it_has_no_counterpart():
    text = open(os.path.join(ROOT, "include", "mrgpu.h")).read()
    # the comment block preceding each prototype names a reference file or states that there is none
    for m in re.finditer(r"/\*((?:(?!\*/).)*?)\*/\s*(?:int|const char \*)\s*(mrgpu_[a-z0-9_]+)\s*\(", text, flags=re.S):
        comment, name = m.group(1), m.group(2)
        assert re.search(r"(src|oink|cpu|examples)/[A-Za-z_0-9]+\.(cpp|h)|[Nn]o counterpart|MapReduce::|KeyValue::|ncclGetUniqueId|ncclCommInitRank|message of the last", comment), name


def test_no_cpu_fallback_without_a_device():
    import torch

    if torch.cuda.is_available():
        return
    try:
        g.MapReduce()
    except g.MRError as e:
        assert "no CUDA device" in str(e)
    else:
        raise AssertionError("constructing a MapReduce without a GPU must fail")


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "gpu_mapreduce_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dp, f), errors="replace").read()
                assert "mr_oracle" not in src and "oracle/" not in src and "mrhelpers" not in src, f


def test_settings_are_documented_and_known_to_the_library():
    """every setting the Python mirror exposes beside the reference's public members is explained in include/mrgpu.h
    and listed in INTEGRATION.md, and is a name the library's settings table knows (read from the source: no device
    is needed)"""
    ref_members = {"mapstyle", "all2all", "verbosity", "timer", "memsize", "minpage", "maxpage", "freepage", "outofcore",
                   "zeropage", "keyalign", "valuealign"}
    extra = [s for s in g.MapReduce._SETTINGS if s not in ref_members and s != "debug_fail_alloc"]
    assert {"compact", "ahead", "combine", "presorted", "l2pin", "hbm_budget"} <= set(extra)
    header = open(os.path.join(ROOT, "include", "mrgpu.h")).read()
    integ = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    table = open(os.path.join(ROOT, "gpu_mapreduce_b200", "csrc", "mrgpu.cu")).read()
    for s in g.MapReduce._SETTINGS:
        assert f'{{"{s}", &mr->{s}}}' in table.replace("  ", " "), s
    for s in extra:
        assert f'"{s}"' in header, s
        assert f"`{s}`" in integ, s
