"""GPU test of mrgpu_sort_pairs_noalloc, the entry point for callers that cannot allocate GPU memory
(an application whose arena already owns the device's memory, or a latency-critical path that cannot absorb
cudaMalloc's device-wide synchronisation): its scratch is a static __device__ array inside libmrgpu.so.

The file name sorts last on purpose: if this addition fails, a run that stops at the first failure has
already executed every other test.  The test is self-contained (fresh process or after the others).

It calls the new entry point through the C-ABI twice in a row -- the second call re-uses the scratch the first
one dirtied, with another size and other types -- and requires output identical to mrgpu_sort_pairs (integer
data: exact equality, the tolerance the other tests use)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_PAIRS = 1 << 20


def _run(fn_name, keys, pay, kbytes, pbytes):
    import torch

    import gpu_mapreduce_b200 as g

    tk = torch.from_numpy(keys.view(np.int32 if kbytes == 4 else np.int64).copy()).cuda()
    tp = torch.from_numpy(pay.view(np.int32 if pbytes == 4 else np.int64).copy()).cuda() if pbytes else None
    rc = getattr(g.lib(), fn_name)(0, None, tk.data_ptr(), tp.data_ptr() if pbytes else None, len(keys), kbytes, pbytes)
    assert rc == 0, g.lib().mrgpu_last_error(None)
    return tk.cpu().numpy().view(keys.dtype), (tp.cpu().numpy().view(pay.dtype) if pbytes else None)


def _case(seed, n, kbytes, pbytes):
    r = np.random.RandomState(seed)
    kd = np.uint32 if kbytes == 4 else np.uint64
    keys = r.randint(0, 1 << 30, size=n).astype(kd)
    keys[::5] = r.randint(0, 4, size=len(keys[::5]))          # long runs of equal keys: stability matters
    if kbytes == 8:
        keys[::3] <<= np.uint64(31)
    pay = None
    if pbytes:
        pay = (np.arange(n) * 7 + 3).astype(np.uint32 if pbytes == 4 else np.uint64)
    return keys, pay


def test_noalloc_sort_matches_allocating_sort_two_calls_in_a_row():
    # call 1 fills the whole scratch (maximum size, widest types); call 2 is smaller and narrower, so stale
    # histogram / look-back / ticket state from call 1 would show up as a wrong order
    for seed, n, kb, pb in [(1, MAX_PAIRS, 8, 8), (2, 300_001, 4, 4)]:
        keys, pay = _case(seed, n, kb, pb)
        want_k, want_p = _run("mrgpu_sort_pairs", keys, pay, kb, pb)
        got_k, got_p = _run("mrgpu_sort_pairs_noalloc", keys, pay, kb, pb)
        assert np.array_equal(got_k, want_k)
        assert np.array_equal(got_p, want_p)
        order = np.argsort(keys, kind="stable")                # and both are the stable sort
        assert np.array_equal(got_k, keys[order]) and np.array_equal(got_p, pay[order])


def test_noalloc_sort_keys_only_and_empty():
    keys, _ = _case(3, 77_777, 4, 0)
    want_k, _ = _run("mrgpu_sort_pairs", keys, None, 4, 0)
    got_k, _ = _run("mrgpu_sort_pairs_noalloc", keys, None, 4, 0)
    assert np.array_equal(got_k, want_k)
    import gpu_mapreduce_b200 as g

    assert g.lib().mrgpu_sort_pairs_noalloc(0, None, None, None, 0, 4, 0) == 0


def test_noalloc_sort_rejects_sizes_beyond_its_scratch():
    import torch

    import gpu_mapreduce_b200 as g

    n = MAX_PAIRS + 1
    tk = torch.zeros(n, dtype=torch.int32, device="cuda")
    rc = g.lib().mrgpu_sort_pairs_noalloc(0, None, tk.data_ptr(), None, n, 4, 0)
    assert rc == -5                                            # MRGPU_ELIMIT
    msg = g.lib().mrgpu_last_error(None).decode()
    assert "1048576" in msg and "at most" in msg
