"""OINK, the reference's script front-end (oink/*.cpp), compiled UNMODIFIED twice by the build step in the container:
against the reference library (oracle/_ref/oink_ref) and against this repo's mirror classes over libmrgpu.so with
cpp/oink_gpu_register.cpp tying rmat_generate / cull / count / edge_to_vertices ... to their device functors
(tests/cpp/_bin/oink_gpu).  Both run the same input scripts; messages and output files must agree.
SURVEY.md B.5 is the known answer for the first script."""
import os
import re
import subprocess
import tempfile

import pytest

from gpu_mapreduce_b200 import workloads as W

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OINK_REF = os.path.join(ROOT, "oracle", "_ref", "oink_ref")
OINK_GPU = os.path.join(ROOT, "tests", "cpp", "_bin", "oink_gpu")

needs_ref = pytest.mark.skipif(not os.path.exists(OINK_REF), reason="reference OINK not built (run __graft_entry__.build() in the container)")
needs_both = pytest.mark.skipif(not (os.path.exists(OINK_REF) and os.path.exists(OINK_GPU)), reason="OINK binaries not built")

RMAT_DEGREE = """rmat 12 8 0.57 0.19 0.19 0.05 0.0 12345 -o tmp.rmat mre
degree 2 -i mre -o tmp.degree mrd
degree 1 -i mre -o tmp.degree1 NULL
degree_stats 2 -i mre
histo -i mrd -o NULL mrh
"""
GRAPH = """rmat 9 4 0.45 0.25 0.15 0.15 0.0 777 -o NULL mre
edge_upper -i mre -o tmp.upper mre
vertex_extract -i mre -o tmp.vert NULL
neighbor -i mre -o tmp.neigh NULL
tri_find -i mre -o tmp.tri mrt
"""
CC = """rmat 9 2 0.25 0.25 0.25 0.25 0.0 4242 -o NULL mre
edge_upper -i mre -o NULL mre
cc_find 0 -i mre -o tmp.cc mrc
"""


def canon_line(line):
    """Canonical form of an output line: first token (the key), then the remaining tokens sorted.  The list a key carries
    (a vertex's neighbours) is in value ARRIVAL order, which follows the order of the pairs in the KeyValue it was built
    from; the GPU library emits KMV pairs in key order where the reference emits them in first-appearance order, so such
    lists are compared after sorting, as BASELINE.json's north star allows ("compared after canonical sorting")."""
    t = line.split()
    return " ".join(t[:1] + sorted(t[1:]))


def run_oink(binary, script, files=()):
    """run `binary < script` in a fresh directory; returns (messages, {output file: sorted lines})"""
    with tempfile.TemporaryDirectory() as td:
        for name, data in files:
            with open(os.path.join(td, name), "wb") as f:
                f.write(data)
        p = subprocess.run([binary], input=script, capture_output=True, text=True, cwd=td, timeout=900)
        assert p.returncode == 0, (p.stdout[-2000:], p.stderr[-2000:])
        msgs = [ln for ln in p.stdout.splitlines() if not re.search(r"secs|NCCL version", ln)]
        outs = {}
        for fn in sorted(os.listdir(td)):
            if fn.startswith("tmp."):
                with open(os.path.join(td, fn)) as f:
                    outs[fn] = sorted(canon_line(ln) for ln in f.read().splitlines())
        return msgs, outs


@needs_ref
def test_reference_oink_known_answer():
    msgs, outs = run_oink(OINK_REF, RMAT_DEGREE)
    assert "RMAT: 4096 rows, 32768 non-zeroes, 7 iterations" in msgs
    assert "Degree: 3055 vertices, 32768 edges" in msgs
    assert len(outs["tmp.rmat.0"]) == 32768 and len(outs["tmp.degree.0"]) == 3055


@pytest.mark.gpu
@needs_both
@pytest.mark.parametrize("script", [RMAT_DEGREE, GRAPH, CC], ids=["rmat_degree", "graph", "cc"])
def test_oink_scripts_same_results_on_both_libraries(script):
    ref_msgs, ref_out = run_oink(OINK_REF, script)
    gpu_msgs, gpu_out = run_oink(OINK_GPU, script)
    assert gpu_msgs == ref_msgs
    assert sorted(gpu_out) == sorted(ref_out)
    for fn in ref_out:
        assert gpu_out[fn] == ref_out[fn], fn


@pytest.mark.gpu
@needs_both
def test_oink_wordfreq_command():
    text = W.words_text(40000, nvocab=300, exponent=1.2)
    script = "wordfreq 10 -i words.txt -o tmp.wf NULL\n"
    ref_msgs, ref_out = run_oink(OINK_REF, script, [("words.txt", text)])
    gpu_msgs, gpu_out = run_oink(OINK_GPU, script, [("words.txt", text)])
    # the fork's library prints "nproc and #map task is ..." from its file map (src/mapreduce.cpp:1084-1085): not OINK output;
    # the top-10 list orders words of equal count by an unspecified qsort order: compare its counts, and the whole table by file
    def clean(msgs):
        return [m for m in msgs if not m.startswith("nproc and #map task")]

    ref_msgs, gpu_msgs = clean(ref_msgs), clean(gpu_msgs)
    assert gpu_msgs[-1] == ref_msgs[-1] == "WordFreq: 1 files, 40000 words, 300 unique"
    assert [m.split()[0] for m in gpu_msgs[1:-1]] == [m.split()[0] for m in ref_msgs[1:-1]] and len(ref_msgs) == 12
    assert gpu_out == ref_out and len(ref_out["tmp.wf.0"]) == 300
