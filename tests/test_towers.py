"""Tower plumbing without GPUs (BASELINE config 1): 2 CPU towers over gloo, SmallResNet, 32x32x3, batch 16, 10 classes."""
import os
import socket
import subprocess
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(tmp, mode, world):
    env = dict(os.environ, OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), os.path.join(HERE, "_tower_worker.py"),
           str(tmp), mode]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return [torch.load(os.path.join(tmp, f"{mode}_w{world}_r{k}.pt")) for k in range(world)]


@pytest.fixture(scope="module")
def runs(tmp_path_factory):
    tmp = tmp_path_factory.mktemp("towers")
    return {("frozen_bn", 2): _run(tmp, "frozen_bn", 2), ("frozen_bn", 1): _run(tmp, "frozen_bn", 1),
            ("train_bn", 2): _run(tmp, "train_bn", 2)}


def test_mirrors_bit_identical_after_every_step(runs):
    for key in (("frozen_bn", 2), ("train_bn", 2)):
        a, b = runs[key]
        assert a["step"] == b["step"] == 3
        for pa, pb in zip(a["per_step"], b["per_step"]):
            assert torch.equal(pa, pb)


def test_two_towers_equal_one_tower_on_the_same_global_batch(runs):
    """Mean of per-tower mean-loss gradients == gradient of the global mean loss (BN frozen)."""
    two, one = runs[("frozen_bn", 2)][0], runs[("frozen_bn", 1)][0]
    assert torch.allclose(two["params"], one["params"], atol=2e-5), float((two["params"] - one["params"]).abs().max())


def test_moving_statistics_are_the_tower_mean(runs):
    a, b = runs[("train_bn", 2)]
    assert torch.equal(a["stats"], b["stats"])          # synced (X3) ...
    frozen = runs[("frozen_bn", 2)][0]["stats"]
    assert not torch.equal(a["stats"], frozen)           # ... and actually updated by training-mode BN


def test_logged_loss_is_tower_mean_and_metric_sums_are_global(tmp_path):
    """X5 / X4 (model.py:491-500, core/metric.py:42): every tower logs the MEAN of the towers' losses, and the metric
    sums that feed the streaming means are the tower mean of the per-tower sums (x world = the global sum)."""
    a, b = _run(tmp_path, "trainer_scalars", 2)
    for ra, rb in zip(a, b):
        assert ra["local"] != rb["local"]                                   # different data per tower ...
        assert ra["logged"] == rb["logged"]                                 # ... one logged value everywhere
        assert abs(ra["logged"] - 0.5 * (ra["local"] + rb["local"])) < 1e-6
        want = 0.5 * (ra["top1_local_sum"] + rb["top1_local_sum"])
        assert abs(ra["top1_mean_of_sums"] - want) < 1e-6 and ra["top1_mean_of_sums"] == rb["top1_mean_of_sums"]


def test_a_dead_tower_is_detected_not_waited_for(tmp_path):
    """Fault injection: tower 1 exits abruptly before step 1; tower 0 must fail with an error from the
    collective (peer closed / timeout) instead of hanging, and the launcher must report a non-zero exit."""
    env = dict(os.environ, OMP_NUM_THREADS="2", TDL_COLLECTIVE_TIMEOUT_S="15")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(HERE, "_tower_worker.py"), str(tmp_path),
           "rank1_dies"]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode != 0
    assert not os.path.exists(os.path.join(tmp_path, "rank1_dies_w2_r0.pt"))  # tower 0 never reported success


def test_descriptor_exchange_between_towers(tmp_path):
    """The VMM symmetric heap shares its allocations as POSIX descriptors over UNIX sockets (parallel/p2p.py): the
    all-gather form delivers every tower's descriptor to every tower, the broadcast form only tower 0's."""
    for rec in _run(tmp_path, "fd_exchange", 3):
        assert rec["texts"] == ["tower 0", "tower 1", "tower 2"]
        assert rec["root_text"] == "tower 0" and rec["others_empty"]

