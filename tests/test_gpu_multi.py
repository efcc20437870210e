"""Two ranks on two B200s through the drop-in command lines (torchrun): the sharded Philox solve and
its single NCCL all-reduce must reproduce the single-process result bit for bit (integer tallies,
same stream ids).  Skipped on a one-GPU box; the CPU suite covers the same host logic over gloo
(tests/test_dist_gloo.py)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import PKG

pytestmark = pytest.mark.gpu


def _two_gpus():
    import torch
    return torch.cuda.is_available() and torch.cuda.device_count() >= 2


def _torchrun(script, argv, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(PKG, script)] + argv
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if r.returncode != 0:
        tb = r.stderr.find("Traceback")
        raise AssertionError(r.stderr[tb:tb + 3000] if tb >= 0 else r.stderr[-3000:])


@pytest.mark.skipif(not _two_gpus(), reason="needs 2 GPUs")
def test_fluence_cli_two_ranks_equals_one(tmp_path):
    import fluence_mini_app as app
    common = ["-r", "pc", "--rng", "philox", "-M", "1001", "-nt", "64", "--rand_seed", "3", "-sa", "0.4"]
    two = ["--matrix_filename", str(tmp_path / "m2.csv"), "--results_filename", str(tmp_path / "r2.csv"),
           "--log_file", str(tmp_path / "log2" / "x.log")]
    _torchrun("fluence_mini_app.py", common + two, 29631)
    a = vars(app.build_parser().parse_args(common))
    a.update(matrix_filename=str(tmp_path / "m1.csv"), results_filename=str(tmp_path / "r1.csv"))
    np.random.seed(3)
    one = app.run(a)
    got = np.genfromtxt(str(tmp_path / "r2.csv"), delimiter=",")
    assert np.array_equal(got, np.genfromtxt(str(tmp_path / "r1.csv"), delimiter=","))
    assert np.array_equal(got, one) and got.any()


@pytest.mark.skipif(not _two_gpus(), reason="needs 2 GPUs")
def test_simple_flux_cli_two_ranks_equals_one(tmp_path):
    import simple_flux_miniapp as app
    common = ["-r", "pc", "--rng", "philox", "-M", "200001", "--rand_seed", "5", "--use_sinks"]
    two = ["--matrix_filename", str(tmp_path / "m2.csv"), "--results_filename", str(tmp_path / "r2.csv"),
           "--log_file", str(tmp_path / "log2" / "x.log")]
    _torchrun("simple_flux_miniapp.py", common + two, 29632)
    a = vars(app.build_parser().parse_args(common))
    a.update(matrix_filename=str(tmp_path / "m1.csv"), results_filename=str(tmp_path / "r1.csv"))
    np.random.seed(5)
    one = app.run(a)
    got = np.genfromtxt(str(tmp_path / "r2.csv"), delimiter=",")
    assert got.shape == (2, 1002) and np.array_equal(got, one)


@pytest.mark.skipif(not _two_gpus(), reason="needs 2 GPUs")
@pytest.mark.parametrize("p2p,n", [("1", "10000"), ("1", "4097"), ("0", "10000")])
def test_p2p_score_allreduce_equals_nccl(p2p, n):
    """nrw.dist.ScoreAllReduce: the one-kernel reduction over NVLink peer memory (and its NCCL
    fall-back) equals dist.all_reduce over 25 back-to-back epochs on both score buffers."""
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "p2p_reduce_worker.py")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(29641 + int(n) % 7 + (0 if p2p == "1" else 10)), worker]
    env = dict(os.environ, NRW_P2P_REDUCE=p2p, NRW_TEST_N=n)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "p2p_reduce_worker ok p2p=%s" % (p2p == "1") in r.stdout, (r.stdout[-1500:], r.stderr[-3000:])


@pytest.mark.skipif(not _two_gpus(), reason="needs 2 GPUs")
@pytest.mark.parametrize("p2p", ["1", "0"])
def test_fused_walk_exchange_two_ranks(p2p):
    """The walk's tail sums over ranks (nrw_walk.reduce): equal to dist.all_reduce and to the unsharded walk, through
    TallyExchange, solve_fluence and the host entry; p2p = 0: the logged NCCL fall-back."""
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fused_exchange_worker.py")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(29671 + int(p2p)), worker]
    env = dict(os.environ, NRW_P2P_REDUCE=p2p)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=400, env=env)
    if r.returncode != 0:
        print(r.stderr[-8000:])
    assert r.returncode == 0 and "fused_exchange_worker ok p2p=%s" % (p2p == "1") in r.stdout, (r.stdout[-1500:], r.stderr[-3000:])
