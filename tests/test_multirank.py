"""N > 1 path.  CPU: two gloo ranks exercise the host-side exchange plan of aggregate() (mrgpu_plan_exchange)
against the reference's delivery rule.  GPU (needs >= 2 devices): the real aggregate/collate over NCCL."""
import os
import signal
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "multi_rank_worker.py")


def _launch(nproc, backend, port, timeout=240, **extra):
    env = dict(os.environ, MR_WORKER_BACKEND=backend, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1", **extra)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), WORKER]
    # own process group: a rank stuck in a collective must not outlive the test
    p = subprocess.Popen(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, start_new_session=True)
    try:
        out, err = p.communicate(timeout=timeout)
    except subprocess.TimeoutExpired:
        os.killpg(p.pid, signal.SIGKILL)
        out, err = p.communicate()
        raise AssertionError(("timed out", out[-2000:], err[-4000:]))
    if not (p.returncode == 0 and "MULTI_OK" in out):
        # pytest shortens long assertion messages: keep the whole output where a gpurun call brings it back
        d = os.path.join(ROOT, "gpurun_out")
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, f"multirank_fail_{backend}_{nproc}_{port}.log"), "w") as f:
            f.write(out + "\n---- stderr ----\n" + err)
    assert p.returncode == 0 and "MULTI_OK" in out, (out[-2000:], err[-4000:])


def test_exchange_plan_two_gloo_ranks():
    _launch(2, "gloo", 29611)


def test_exchange_plan_three_gloo_ranks():
    _launch(3, "gloo", 29612)


@pytest.mark.gpu
def test_aggregate_collate_over_nccl():
    import torch

    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs (run with gpurun --gpus 2)")
    only = os.environ.get("MR_RANKS")            # e.g. MR_RANKS=8: just that rank count (a large box is charged per GPU)
    if only:
        _launch(int(only), "nccl", 29618, timeout=600)
        _launch(int(only), "nccl", 29619, timeout=600, MRGPU_P2P="0")
        return
    _launch(2, "nccl", 29613)                   # exchange by peer stores (CUDA IPC) where the pairs are uniform
    _launch(2, "nccl", 29615, MRGPU_P2P="0")    # the same through ncclSend/ncclRecv
    _launch(2, "nccl", 29616, MRGPU_PEER_DIRECT="1")
    if n >= 3:
        _launch(3, "nccl", 29614)
    if n >= 8:
        _launch(8, "nccl", 29617, timeout=600)
