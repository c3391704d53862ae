"""Two ranks on two GPUs (skipped on a one-GPU box): the distributed vector's device ghost exchange -- update_ghost_values /
compress(add) as the fused NVLink peer-store kernel (default), over NCCL (MGB_P2P=0) and as the send/recv kernel pair
(MGB_P2P_FUSED=0), the split cell loop (MGB_OVERLAP=1), the
V-cycle replayed from a CUDA graph with the exchanges captured, and the chunked cell loop's ghost path (MGB_CHUNKED=1) --
against the serial oracle: vmult <= 1e-12, V-cycle <= 1e-11, identical CG iteration counts (tools/check_multigpu.py, one
process per GPU under torch.distributed.run; reference: LinearAlgebra::distributed::Vector as used in
multigrid_throughput.cc:1143-1144 and include/operator.h:152-183)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _two_gpus():
    import torch
    return torch.cuda.is_available() and torch.cuda.device_count() >= 2


def _run(extra_env, only, port):
    env = dict(os.environ)
    env.update(extra_env)
    if only:
        env["MGB_CHECK_ONLY"] = only
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tools", "check_multigpu.py")]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    bad = [l for l in (r.stdout + "\n" + r.stderr).splitlines() if any(k in l for k in ("FAIL", "Error", "error", "Traceback", "trap"))]
    assert r.returncode == 0 and "MULTIGPU PARITY OK" in r.stdout, "\n".join(bad[:12])
    # every rank must hold cells in every case (a case where one rank owns nothing proves nothing about the exchange)
    for line in r.stdout.splitlines():
        if line.startswith("[rank") and "owned=" in line:
            assert "owned=0 " not in line, line
    return r.stdout


@pytest.mark.skipif(not _two_gpus(), reason="needs two GPUs")
def test_two_ranks_nccl_all_hierarchies():
    _run({}, None, 29541)


@pytest.mark.skipif(not _two_gpus(), reason="needs two GPUs")
@pytest.mark.parametrize("name,env", [("nccl", {"MGB_P2P": "0"}), ("p2p_two_kernels", {"MGB_P2P": "1", "MGB_P2P_FUSED": "0"}),
                                      ("overlap", {"MGB_OVERLAP": "1"}), ("overlap_nccl", {"MGB_OVERLAP": "1", "MGB_P2P": "0"}),
                                      ("no_graph", {"MGB_GRAPH_MULTIRANK": "0"}), ("chunked", {"MGB_CHUNKED": "1"})])
def test_two_ranks_exchange_variants(name, env):
    _run(env, "DefaultPolicy", 29542 + sum(map(ord, name)) % 50)
