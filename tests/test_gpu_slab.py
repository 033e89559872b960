"""GPU: the y-slab driver (pack/unpack kernels, stage ABI, halo path) with world = 1 must equal the fused
single-GPU pipeline bit for bit; with >= 2 GPUs the same check runs under torchrun with NCCL."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("path", ["peer", "nccl"])
@pytest.mark.parametrize("dims", [(32, 32, 16), (32, 256, 8)], ids=["staged_fft", "col_fft_chunks"])
def test_slab_world1_bit_identical(cuda_lib, dims, path, monkeypatch):
    """dims (32, 256, 8): the y passes run on the lanes-along-the-lines kernel in its chunk-layout (SPLIT) form.
    path peer: the exchange arena, the halo push kernel, the device barrier and the transforms that write the peers' receive
    buffers, with this rank as its own neighbour; path nccl: the pack / unpack kernels around torch.distributed."""
    import torch
    from tests.slab_parity import run
    monkeypatch.setenv("MHH_SLAB_PATH", path)
    out, status, ref = run(0, 1, torch.device("cuda:0"), dims=dims)
    for k in ("u", "v", "w", "th", "p"):
        assert np.array_equal(out[k], ref[k]), k
    assert status["dt"] == ref["status"]["dt"]


@pytest.mark.parametrize("mode", ["fused", "fused_pipe", "ce", "ce_overlap", "ce_sub"])
@pytest.mark.parametrize("dims", [(64, 64, 32), (32, 256, 8)], ids=["staged_fft", "col_fft_chunks"])
def test_slab_world1_all_to_all_modes_bit_identical(cuda_lib, dims, mode, monkeypatch):
    """Every way the peer path moves the all-to-all chunks (MHH_SLAB_A2A) with this rank as its own peer; ce_sub runs the
    solve on 4 kx sub-blocks (segment tables, sub-block y transforms, Thomas sweep with the sub-block's kx offset)."""
    import torch
    from tests.slab_parity import run
    monkeypatch.setenv("MHH_SLAB_PATH", "peer")
    monkeypatch.setenv("MHH_SLAB_A2A", mode)
    out, status, ref = run(0, 1, torch.device("cuda:0"), dims=dims)
    for k in ("u", "v", "w", "th", "p"):
        assert np.array_equal(out[k], ref[k]), k
    assert status["dt"] == ref["status"]["dt"]


@pytest.mark.parametrize("path", ["peer", "nccl"])
def test_slab_world1_order4_bit_identical(cuda_lib, path, monkeypatch):
    """The 4th-order pipeline (advec_4 + diff_4, Pres_4 with its 7-band solve, three-deep halos, uflux forcing) through the
    slab driver's stage calls equals the single-GPU pipeline bit for bit."""
    import torch
    from tests.slab_parity import run
    monkeypatch.setenv("MHH_SLAB_PATH", path)
    out, status, ref = run(0, 1, torch.device("cuda:0"), dims=(32, 32, 24), case="dns4")
    for k in ("u", "v", "w", "p"):
        assert np.array_equal(out[k], ref[k]), k
    assert status["dt"] == ref["status"]["dt"]


@pytest.mark.parametrize("uflux", ["0", "1"], ids=["bitwise", "uflux_1e-13"])
@pytest.mark.parametrize("path", ["peer", "nccl"])
def test_slab_two_ranks_order4(cuda_lib, path, uflux):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29534", os.path.join(ROOT, "tests", "slab_parity.py")], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, SLAB_DIMS="64,64,32", MHH_SLAB_PATH=path, SLAB_CASE="dns4", SLAB_UFLUX=uflux))
    print(r.stdout[-3000:], r.stderr[-2000:])
    assert r.returncode == 0 and "SLAB PARITY OK" in r.stdout


@pytest.mark.parametrize("mode", ["fused", "ce_sub"])
def test_slab_two_ranks_all_to_all_modes(cuda_lib, mode):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29535", os.path.join(ROOT, "tests", "slab_parity.py")], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, SLAB_DIMS="64,512,16", MHH_SLAB_PATH="peer", MHH_SLAB_A2A=mode))
    print(r.stdout[-3000:], r.stderr[-2000:])
    assert r.returncode == 0 and "SLAB PARITY OK" in r.stdout


@pytest.mark.parametrize("path", ["peer", "nccl"])
@pytest.mark.parametrize("dims", ["64,64,32", "32,512,8"], ids=["staged_fft", "col_fft_chunks"])
def test_slab_two_ranks_nccl(cuda_lib, dims, path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29533", os.path.join(ROOT, "tests", "slab_parity.py")], capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, SLAB_DIMS=dims, MHH_SLAB_PATH=path))
    print(r.stdout[-3000:], r.stderr[-2000:])
    assert r.returncode == 0 and "SLAB PARITY OK" in r.stdout
