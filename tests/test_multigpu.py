"""T3: fused push/pull across GPUs (needs >= 2 devices; each case is a torchrun job)."""
import os
import socket
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _run(n, *args, timeout=180):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "multigpu_worker.py")] + list(args)
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout, cwd=ROOT)
    assert r.returncode == 0 and "MULTIGPU_OK" in r.stdout, r.stdout[-3000:]


@pytest.mark.parametrize("model", ["lstm", "encoder", "gan"])
def test_sync_push_matches_oracle_and_is_deterministic(model):
    _run(2, "--model", model, "--mode", "sync")


def test_sync_push_big_tile_path_across_ranks(monkeypatch):
    """The 128x128-tile path of the push kernel (16-byte stores into every rank's arena), forced at the
    encoder's sizes: same oracle, identical-arena and determinism checks as the small-tile path."""
    monkeypatch.setenv("PSB200_BIG_TILES_MIN", "1024")
    _run(2, "--model", "encoder", "--mode", "sync")


@pytest.mark.parametrize("model", ["lstm", "gan"])
def test_async_push_counts_every_update(model):
    _run(2, "--model", model, "--mode", "async")


@pytest.mark.parametrize("model", ["lstm", "encoder", "gan"])
def test_pull_fused_into_consumer_kernels(model):
    """bcast=False: compute copies stay in the owner's arena; the LSTM kernels and the dense GEMMs
    (psb_gemm_bf16_pull) read them through peer pointers."""
    _run(2, "--model", model, "--mode", "sync", "--pull_from_owner", "1")


@pytest.mark.parametrize("model", ["lstm"])
def test_signal_protocol_under_host_jitter(model):
    """T4 stress: random per-rank host stalls and a mix of blocking / pipelined submissions must not
    change a single bit of the sync-mode result (device-side step tokens order everything)."""
    _run(2, "--model", model, "--mode", "sync", "--stress", "60")


@pytest.mark.parametrize("model", ["lstm", "encoder"])
def test_sync_r_of_n_keeps_stepping_past_a_slow_rank(model):
    """--replicas_to_aggregate < workers on the NVLink transport: ps service thread per
    rank, first-R aggregation, stale gradients dropped, tokens."""
    _run(2, "--model", model, "--mode", "sync", "--replicas", "1", "--slow_rank", "1", "--steps", "40")


@pytest.mark.parametrize("model", ["lstm", "gan"])
def test_async_owner_side_apply_counts_every_update(model):
    _run(2, "--model", model, "--mode", "async", "--async_apply", "owner", "--steps", "20")


def test_stalled_peer_raises_instead_of_hanging():
    _run(2, "--model", "lstm", "--mode", "sync", "--stall", "1")


def test_noalloc_entry_point_across_ranks():
    _run(2, "--model", "lstm", "--mode", "sync", "--noalloc", "1")


def test_nvl_entrypoints_through_the_reference_cli(tmp_path):
    """examples/distributed_lstm.py --backend=nvl: 1 ps (done-queue only) + 2 GPU workers, sync."""
    from psb200.launch import launch
    rc, logs = launch("lstm", workers=2, ps=1, backend="nvl",
                      extra=["--synthetic", "--sync", "--num_steps=40", "--save_model_secs=1"],
                      model_dir=str(tmp_path / "m"), quiet=True, timeout=240)
    assert rc == 0, logs
    assert "ps 0 received done 1" in logs["ps:0"] and "ps 0: quitting" in logs["ps:0"]
    for w in ("worker:0", "worker:1"):
        assert "Global step 40 Local step 39" in logs[w], logs[w][-2000:]
        assert "Test-Accuracy:" in logs[w]
    assert any(f.startswith("model.ckpt-") for f in os.listdir(tmp_path / "m"))
