"""The opt-in peer-memory gradient exchange (K4c: reduce-scatter + SGD + all-gather of a bucket in one kernel over
NVLink, csrc/peer_exchange.cu) on two GPUs of one box; skipped with fewer.  Checker at kernel level: the closed form
(sum over ranks in rank order, then the update of b200_sgd_flat_f32, pinned by tests/test_gpu_kernels.py); at model
level: the default FlatDDP(exchange="nccl") path, itself pinned against stock DistributedDataParallel."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_gpu_peer_exchange_matches_closed_form_and_nccl_path():
    env = dict(os.environ, PYTHONPATH=ROOT, B200_PEER_TIMING="0")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29593", os.path.join(ROOT, "tests", "peer_exchange_worker.py")]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    k = out["kernel"]
    assert k["error_word"] == 0 and k["step_counter"] == 3          # no flag timed out; three iterations counted
    assert k["replicas_identical"] and k["shadow_ok"]              # every rank holds the same bits, shadow = bf16(p)
    assert k["max_err"] <= k["bound"], k                           # float32, FMA contraction only
    m = out["model"]
    assert not m["peer"]["error"] and m["peer"]["identical"] and m["nccl"]["identical"]
    assert m["peer"]["steps"] == m["nccl"]["steps"] == 3
    for a, b in zip(m["peer"]["epoch"], m["nccl"]["epoch"]):
        assert a == pytest.approx(b, rel=2e-4, abs=1e-6)           # same losses: only the summation order differs
    assert m["global_rel_weight_diff"] < 1e-4, m   # measured 1.5e-5: summation order + fused update
    # the same exchange with the whole iteration captured as one CUDA graph (bf16 activations, 8 iterations)
    g = out["graph"]
    assert g["peer"]["captured"] and not g["peer"]["error"] and g["peer"]["steps"] == g["nccl"]["steps"] == 8
    assert g["peer"]["epoch"][0] == pytest.approx(g["nccl"]["epoch"][0], rel=2e-3) and g["peer"]["epoch"][1:] == g["nccl"]["epoch"][1:]
    assert g["peer"]["w"] == pytest.approx(g["nccl"]["w"], rel=1e-4) and g["peer"]["w1"] == pytest.approx(g["nccl"]["w1"], rel=1e-4)
    # BatchNorm buffers: rank 0's running statistics reach every rank through the peer copy, as through NCCL
    assert g["peer"]["rv"] == pytest.approx(g["nccl"]["rv"], rel=1e-3)
    assert g["peer"]["rv_all"] and g["nccl"]["rv_all"]
    # train(..., graph=True) on the peer path: the public switch picks the single-graph runner and gives the same run
    a = g["peer_auto"]
    assert not a["error"] and a["steps"] == 8 and a["rv_all"]
    assert a["epoch"][0] == pytest.approx(g["nccl"]["epoch"][0], rel=2e-3) and a["epoch"][1:] == g["nccl"]["epoch"][1:]
    assert a["w"] == pytest.approx(g["nccl"]["w"], rel=1e-4) and a["rv"] == pytest.approx(g["nccl"]["rv"], rel=1e-3)
    assert g["segmented_refuses_peer"]


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_gpu_peer_exchange_gives_up_cleanly_when_a_peer_never_arrives():
    """~20 s: rank 1 never launches its half of the exchange.  Rank 0 must time out, report it, and nobody's parameters,
    shadow or momentum may have been written (no update from gradients that were never known to be final)."""
    env = dict(os.environ, PYTHONPATH=ROOT, B200_PEER_TIMING="0", B200_PEER_MODEL="0", B200_PEER_GRAPH="0",
               B200_PEER_TIMEOUT="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29594", os.path.join(ROOT, "tests", "peer_exchange_worker.py")]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    t = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])["timeout"]
    assert t["error_word_rank0"] == 1 and t["step_counter"] == 0
    assert t["all_arenas_untouched_on_all_ranks"]
    assert 5.0 < t["waited_s"] < 120.0 and t["second_attempt_s"] < 2.0
