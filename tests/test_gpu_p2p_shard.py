"""Output-channel sharding with the library's own exchange over peer memory
(dashb_shard_export / _import / _exchange): `world` processes, one "rank"
each, all on GPU 0, map each other's gather buffers through CUDA IPC and run
the real protocol.  Every rank must end with the reference's output labels."""
import os
import socket
import subprocess
import sys

import pytest

from helpers import ROOT

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.parametrize("name,world,batch,lanes", [("MLP_TINY", 2, 40, 2), ("CONV_UNIT", 3, 33, 2), ("RESCALE_OPT", 2, 5, 1),
                                                    ("SIGN_TINY", 4, 64, 2), ("GARBLED:CIFAR_TINY_SCALE", 3, 40, 2),
                                                    ("GARBLED:MNIST_TINY_CNN", 4, 48, 2)])
@pytest.mark.parametrize("compress", ["1", "0"])
def test_peer_memory_exchange_between_processes(gpu, name, world, batch, lanes, compress):
    """compress = 1: labels cross as 16-byte compressed values and are expanded
    on arrival (k_shard_push_comp / k_shard_unpack, the default); 0: the digit
    planes themselves (k_shard_push)."""
    if compress == "0" and name in ("RESCALE_OPT", "GARBLED:MNIST_TINY_CNN"):
        pytest.skip("plane exchange: covered by the other circuits")
    port = str(_free_port())
    worker = os.path.join(ROOT, "tests", "p2p_shard_worker.py")
    env = dict(os.environ, DASHB_SHARD_COMPRESS=compress)
    procs = [subprocess.Popen([sys.executable, worker, str(r), str(world), port, name, str(batch), str(lanes)], env=env,
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(world)]
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(out)
    for r, (p, out) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and ("rank %d ok" % r) in out, "rank %d:\n%s" % (r, out[-3000:])


def test_cpp_sharded_example(gpu, tmp_path):
    """integration/example_sharded.cpp: the same sharding driven from plain C++
    (fork, C ABI only, IPC handles passed through files): three ranks on GPU 0
    evaluate a GPU-garbled scaled CNN; rank 0 checks the result against the
    unsharded evaluation, label for label, and that every inference decodes."""
    import workloads as W
    from dash_b200 import garble

    exe = os.path.join(ROOT, "integration", "bin", "example_sharded")
    if not os.path.exists(exe):
        pytest.skip("integration/bin/example_sharded not built")
    spec = W.MODELS["CIFAR_TINY_SCALE"]
    plain, n_in = W.plain_layers(spec, weight_seed=4)
    path = str(tmp_path / "m.gcm")
    garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, path, seed=b"cpp-sharded")
    res = subprocess.run([exe, path, "3", "40", "1"], capture_output=True, text=True, timeout=240)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "sharded == unsharded: 3 ranks, batch 40" in res.stdout
