"""CPU: the HBM pool's bookkeeping (best fit, split, coalesce, growth in place, trim, injected failures) and a store's
growth, with the CUDA runtime calls replaced by host stand-ins (tests/cpp/pool_host_test.cu).  nvcc compiles the test
because the headers hold device code; nothing is launched, so it runs without a GPU."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build_and_run(name, ok, *args):
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not (os.path.exists(nvcc) or shutil.which(nvcc)):
        pytest.skip("no nvcc")
    out = os.path.join(ROOT, "tests", "cpp", "_bin", name)
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O1", "-std=c++17", "--extended-lambda", "-w", "-o", out,
                           os.path.join(ROOT, "tests", "cpp", name + ".cu")])
    res = subprocess.run([out] + list(args), capture_output=True, text=True)
    assert res.returncode == 0 and ok in res.stdout, res.stdout + res.stderr


def test_pool_and_store_growth_on_the_host():
    _build_and_run("pool_host_test", "POOL_HOST_OK")


def test_rmat_generator_arithmetic_against_the_c_library():
    """the restated drand48 step, the k-step jump and the integer form of "drand48() < t" (rmat_lcg.cuh) against glibc's
    srand48 / drand48 / seed48 on the host: what makes the device generator bit-exact is checked without a device"""
    _build_and_run("rmat_host_test", "RMAT_HOST_OK")


def test_device_hash_restatement_against_reference_kats(golden):
    """hash.cuh is __host__ __device__ code: the functions the partition and convert kernels call are run on the host
    against the known answers recorded from the reference's src/hash.cpp, at every alignment, and the grouping
    fingerprint's two bodies (word-wise for aligned keys, byte-wise otherwise) against each other"""
    kats = os.path.join(ROOT, "tests", "cpp", "_bin", "hash_kats.txt")
    os.makedirs(os.path.dirname(kats), exist_ok=True)
    with open(kats, "w") as fp:
        for iv, hexkey, h in golden.manifest["hash"]:
            fp.write(f"{iv} {hexkey or '-'} {h}\n")
    _build_and_run("hash_host_test", "HASH_HOST_OK", kats)


def test_merge_path_and_thread_steps_against_std_merge():
    """merge.cuh's split search and per-thread steps, driven on the host the way the two kernels drive them, against
    std::merge: duplicates, empty runs, lengths around tile boundaries (800 cases)"""
    _build_and_run("merge_host_test", "MERGE_HOST_OK")


def test_count_cache_addressing_over_the_whole_key_range():
    """every key below 2^26 -> (slot, tag) -> the same key back (dense_hash.cuh), 1170-1171 keys per slot"""
    _build_and_run("dense_host_test", "DENSE_HOST_OK")


def test_record_layout_arithmetic_against_the_oracle(oracle):
    """kv_rec_size / kmv_rec_size (common.cuh, the functions every kernel uses to place a record) for every alignment
    pair and a grid of sizes, against the oracle's pinned restatement of the reference's layout"""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not (os.path.exists(nvcc) or shutil.which(nvcc)):
        pytest.skip("no nvcc")
    out = os.path.join(ROOT, "tests", "cpp", "_bin", "layout_host_test")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O1", "-std=c++17", "-w", "-o", out,
                           os.path.join(ROOT, "tests", "cpp", "layout_host_test.cu")])
    cases, want = [], []
    aligns = (1, 2, 4, 8, 16, 32)
    sizes = (0, 1, 2, 3, 4, 5, 7, 8, 9, 15, 16, 17, 31, 33, 100, 255, 1 << 20)
    for ka in aligns:
        for va in aligns:
            for kb in sizes:
                for vb in sizes[:12]:
                    cases.append(f"kv {ka} {va} 0 {kb} {vb}")
                    want.append(oracle.lib.orc_kv_record_size(kb, vb, ka, va))
                for nv in (0, 1, 2, 3, 7, 1000, 1 << 26):
                    mvb = nv * 5
                    cases.append(f"kmv {ka} {va} {nv} {kb} {mvb}")
                    want.append(oracle.lib.orc_kmv_record_size(nv, kb, mvb, ka, va))
    res = subprocess.run([out], input="\n".join(cases) + "\n", capture_output=True, text=True)
    assert res.returncode == 0
    lines = res.stdout.split("\n")[:len(cases)]
    assert len(lines) == len(cases)
    for case, line, w in zip(cases, lines, want):
        koff, voff, size = (int(x) for x in line.split())
        assert size == w, (case, line, w)
        assert koff <= voff <= size


def test_href_automaton_against_the_reference_scanner(oracle):
    """href_dfa.cuh -- the transition function the device's per-byte table is built from -- walked over random
    adversarial HTML on the host: the URLs it delimits are the ones the reference's scanner emits"""
    import numpy as np

    import mrhelpers as H

    out = os.path.join(ROOT, "tests", "cpp", "_bin", "href_host_test")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call(["g++", "-O1", "-x", "c++", "-o", out, os.path.join(ROOT, "tests", "cpp", "href_host_test.cu")])
    r = np.random.RandomState(2024)
    toks = [b"<", b"a", b" ", b"h", b"r", b"e", b"f", b"=", b'"', b"x", b"href", b"<a href=\"", b"u", b">", b"\t", b"<a  href = \"",
            b"<ab href=\"", b"<a hrex=\"", b"<<a href=\"", b"hr", b"<a h", b"\n"]
    path = os.path.join(ROOT, "tests", "cpp", "_bin", "href_case.txt")
    total = 0
    for trial in range(60):
        text = b"".join(toks[r.randint(len(toks))] for _ in range(r.randint(1, 6000)))
        with open(path, "wb") as fp:
            fp.write(text)
        res = subprocess.run([out, path], capture_output=True, text=True)
        assert res.returncode == 0
        got = [text[int(a):int(b)] for a, b in (ln.split() for ln in res.stdout.splitlines())]
        want = [k for k, _ in H.kv_decode(oracle.href(text, "f.html"))]
        assert got == want, trial
        total += len(want)
    assert total > 1000


def test_tokenizer_swar_masks_against_a_byte_loop():
    """tok_swar.cuh (whitespace / NUL masks by SWAR compares, word starts and ends as bit operations) on 600 k random
    16-byte vectors with every delimiter, NULs and bytes with the high bit set, against the byte-wise definitions"""
    _build_and_run("tok_host_test", "TOK_HOST_OK")


def test_compact_sort_keys_of_16_byte_keys():
    """compact_key.cuh: sort key <-> key round trip for random varying-byte sets, and the sort keys order the keys as
    128-bit little-endian integers (what lets an iterative job's later rounds be merged instead of sorted)"""
    _build_and_run("compact_host_test", "COMPACT_HOST_OK")


def test_sort_order_transforms_against_the_c_comparisons():
    """sort_order.cuh: the float / double / string transforms of sort_keys compare like <, > and strcmp (the
    reference's compare functions), incl. signed zeros, infinities, denormals and bytes >= 0x80"""
    _build_and_run("sortorder_host_test", "SORTORDER_HOST_OK")
