"""CPU-side contracts: bench.py reference arm, config flags, results file, topology parsing, launcher."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bench_reference_arm_reports_unavailable():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1",
                        "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0
    d = json.loads(p.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and "unavailable" in d


def test_reference_flags_and_config():
    import argparse
    from tadnn.utils import config as cfg
    ap = cfg.add_reference_flags(argparse.ArgumentParser())
    opt = cfg.from_args(ap.parse_args([]))
    # defaults of cifar_example/sgd-torchad_nn-cifar.lua:34-41
    assert (opt.data, opt.seed, opt.learningRate, opt.batchSize, opt.weightDecay, opt.iterations, opt.threads,
            opt.usegpu) == (False, 1, 1e-3, 1, 0, 1, 1, False)
    assert (opt.trsize, opt.tesize) == (4000, 2000)
    opt = cfg.from_args(ap.parse_args(["-data", "-batchSize", "10", "-usegpu", "-learningRate", "0.01"]))
    assert (opt.trsize, opt.tesize, opt.batchSize, opt.usegpu, opt.learningRate) == (32000, 10000, 10, True, 0.01)
    assert "TADNN_PRINT_DIMS" in cfg.ENV_SWITCHES and "TADNN_SYNC_PROTOTYPE" in cfg.ENV_SWITCHES


def test_results_file_has_reference_fields(tmp_path):
    from tadnn.utils.logging import write_results_file
    p = write_results_file(str(tmp_path), train_size=4000, test_size=2000, batch_size=1, accuracy=12.5,
                           learning_rate=1e-3, threads=1, preprocess_s=0.5, train_s=2.0)
    assert os.path.basename(p) == "SgdAuto--Size:4000--Batch:1.txt"
    txt = open(p).read()
    for field in ("Train set Size:", "Test set Size:", "batchSize:", "Accuracy:", "Learning rate:", "Threads:",
                  "Pre-process Time:", "Training Time:"):
        assert field in txt


def test_topology_parsing():
    from tadnn.utils import topo
    txt = "\x1b[4m\tGPU0\tGPU1\tCPU Affinity\x1b[0m\nGPU0\t X \tNV18\t0-23\nGPU1\tNV18\t X \t0-23\n"
    assert topo.nvlink_link_counts(txt) == {"GPU0-GPU1": 18, "GPU1-GPU0": 18}
    assert topo.describe()["gpus"] == (torch.cuda.device_count() if torch.cuda.is_available() else 0)


def test_native_library_builds_and_exports_the_c_abi():
    import ctypes
    from tadnn import _native
    if not _native.available():
        import build_native
        build_native.build()
    lib = _native.lib()
    assert lib.tadnn_version() >= 100
    header = open(os.path.join(ROOT, "include", "tadnn.h")).read()
    import re
    names = set(re.findall(r"\b(tadnn_[a-z0-9_]+)\s*\(", header))
    assert len(names) > 40
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/tadnn.h but not exported"
    # an entry point refuses bad arguments with a message instead of launching
    assert lib.tadnn_gemm_bf16(None, None) != 0
    assert b"null args" in lib.tadnn_last_error()
    del ctypes


def test_cifar_example_runs_on_cpu(tmp_path):
    env = dict(os.environ, MASTER_PORT="29931")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "cifar_dp.py"), "--results-dir", str(tmp_path),
                        "--miniBatch", "500", "-iterations", "1"], capture_output=True, text=True, timeout=600, env=env)
    assert p.returncode == 0, p.stdout[-1500:] + p.stderr[-1500:]
    assert "ConfusionMatrix" in p.stdout
    assert os.path.exists(os.path.join(str(tmp_path), "SgdAuto--Size:4000--Batch:1.txt"))


def test_fake_world_collective_mismatch_times_out_cleanly():
    from tadnn import mpi
    w = mpi.FakeWorld(2)

    def body(h):
        if h.rank() == 0:
            raise RuntimeError("rank 0 died")
        h.barrier()

    try:
        w.run(body, timeout_s=10)
        assert False
    except RuntimeError as e:
        assert "rank 0 died" in str(e)


def test_python_dash_m_tadnn_describes_the_installation():
    import json
    p = subprocess.run([sys.executable, "-m", "tadnn"], capture_output=True, text=True, timeout=300, cwd=ROOT,
                       env=dict(os.environ, TADNN_PRINT_DIMS="1"))
    assert p.returncode == 0, p.stderr[-1500:]
    d = json.loads(p.stdout)
    assert d["native"]["built"] and d["native"]["missing_entry_points"] == [] and d["native"]["entry_points"] > 50
    assert d["env"] == {"TADNN_PRINT_DIMS": "1"} and "gpus" in d["devices"]


def _free_port():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("script,nproc,args,expect", [
    ("mlp_dp_cpu.py", 2, ["--size", "40", "--batchSize", "5"], "4 syncs"),
    ("mlp_mp.py", 2, [], "|fused - dense|"),
    ("mlp_hybrid_cpu.py", 4, ["--tp", "2", "--size", "40", "--batchSize", "5"], "tp 1/2"),
])
def test_multiprocess_examples_run_on_cpu(script, nproc, args, expect):
    """The launcher path of the examples (examples/launch.sh = torchrun on 127.0.0.1), CPU / gloo."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "examples", script)] + args
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
    assert p.returncode == 0, p.stdout[-1500:] + p.stderr[-1500:]
    assert expect in p.stdout, p.stdout[-1500:]
