"""Multi-GPU checks under pytest (SURVEY section 4 "Multi-GPU tests"): one process per GPU over NCCL bootstrap + the P2P
kernels.  Skipped on boxes with fewer than two GPUs (ranks that wait on each other must not share one GPU).

Each case launches ``scripts/p2p_check.py <mode>`` under torchrun and requires every JSON record the script emits to
say ``ok``:
  correctness -- one-shot / two-shot fused all-reduce + SGD-momentum / Adam vs the ordered fp32 reference, ragged tails,
                 mirrors bit-identical, gradients cleared; packed in-place mean (BN statistics / scalars)
  wire        -- the same with gradients crossing NVLink as bf16 (packed and cleared inside the bucket kernel)
  stress      -- 3000 back-to-back small one-/two-shots with changing payloads (signal-pad races)
  train       -- ResNet-50 steps: p2p (eager, graph, overlap) vs the NCCL baseline path, mirrors identical
  resume      -- checkpoint at step 3 with the optimizer slots sharded over the towers, resume, compare with the
                 uninterrupted run (bit-equal on the deterministic optimizer path; rounding-level through the trainer)
  nvls_small  -- buffers from the library's own VMM heap (cuMemCreate + descriptors over UNIX sockets + cuMulticast):
                 peer mappings readable, NVLS fused step vs the fp32 reference (reports "unavailable" without multicast)
  dead_nccl   -- a tower exits mid-run on the NCCL path: the job ends with an error, nobody waits forever
"""
import json
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0
needs2 = pytest.mark.skipif(NGPU < 2, reason=f"needs >= 2 GPUs, this box has {NGPU}")
WORLDS = sorted({2, NGPU} if NGPU > 2 else {2})


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _launch(world, modes, timeout=900, env_extra=None):
    env = dict(os.environ, **(env_extra or {}))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "scripts", "p2p_check.py"), *modes]
    return subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


def _records(stdout):
    out = []
    for line in stdout.splitlines():
        line = line.strip()
        if line.startswith("{") and line.endswith("}"):
            try:
                out.append(json.loads(line))
            except ValueError:
                pass
    return out


@needs2
@pytest.mark.parametrize("world", WORLDS)
@pytest.mark.parametrize("mode", ["correctness", "wire", "stress", "train", "resume", "nvls_small"])
def test_p2p(world, mode):
    r = _launch(world, [mode])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    recs = [x for x in _records(r.stdout) if x.get("event") != "done"]
    assert recs, r.stdout[-2000:]
    bad = [x for x in recs if not x.get("ok", False)]
    assert not bad, bad[:5]
    if mode == "resume":
        kinds = {x.get("kind") for x in recs if x["test"] == "resume_optimizer"}
        assert kinds == {"sgdm", "adam"} and all(x["bit_equal"] for x in recs if x["test"] == "resume_optimizer")
        assert any(x["test"] == "resume_trainer" for x in recs)


@needs2
def test_dead_tower_on_the_nccl_path_ends_the_job():
    r = _launch(2, ["dead_nccl"], timeout=300, env_extra={"TDL_COLLECTIVE_TIMEOUT_S": "20",
                                                          "TORCH_NCCL_ASYNC_ERROR_HANDLING": "1"})
    assert r.returncode != 0
    assert "every step completed" not in r.stdout
    assert "rank 0 finished step 2" not in r.stdout
