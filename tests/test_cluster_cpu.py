"""T1: 1 ps + 2 workers on localhost (BASELINE.json config #1) for every entrypoint, through the
reference command line, plus sync mode and chief restart -> resume."""
import os
import re
import socket
import subprocess
import sys
import time

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ports(n):
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _launch(model, job, idx, ps, workers, model_dir, extra):
    cmd = [sys.executable, os.path.join(ROOT, "examples", "distributed_%s.py" % model),
           "--ps_hosts=" + ",".join(ps), "--worker_hosts=" + ",".join(workers), "--job_name=" + job,
           "--task_index=%d" % idx, "--workers=%d" % len(workers), "--model_dir=" + model_dir,
           "--synthetic"] + extra
    env = dict(os.environ, OMP_NUM_THREADS="2", MKL_NUM_THREADS="2")
    return subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=ROOT, env=env)


def _run_cluster(model, tmp_path, extra, n_ps=1, n_workers=2, timeout=180):
    ports = _ports(n_ps + n_workers)
    ps = ["127.0.0.1:%d" % p for p in ports[:n_ps]]
    workers = ["127.0.0.1:%d" % p for p in ports[n_ps:]]
    model_dir = str(tmp_path / "ckpt")
    procs = [_launch(model, "ps", i, ps, workers, model_dir, extra) for i in range(n_ps)]
    procs += [_launch(model, "worker", i, ps, workers, model_dir, extra) for i in range(n_workers)]
    outs = []
    deadline = time.time() + timeout
    for p in procs:
        try:
            out, _ = p.communicate(timeout=max(1, deadline - time.time()))
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise AssertionError("cluster timed out; partial output:\n" + "\n".join(outs))
        outs.append(out)
        assert p.returncode == 0, out[-3000:]
    return outs[:n_ps], outs[n_ps:], model_dir


def _global_steps(out):
    return [int(m.group(1)) for m in re.finditer(r"Global step (\d+) Local step (\d+)  AvgTime: [\d.]+ms", out)]


@pytest.mark.parametrize("model,steps", [("encoder", 30), ("gan", 40), ("lstm", 30)])
def test_one_ps_two_workers_async(model, steps, tmp_path):
    ps_out, w_out, model_dir = _run_cluster(model, tmp_path, ["--num_steps=%d" % steps, "--save_model_secs=1"])
    # done-queue shutdown order
    lines = [l for l in ps_out[0].splitlines() if l.startswith("ps 0")]
    assert lines == ["ps 0 received done 0", "ps 0 received done 1", "ps 0: quitting"], ps_out[0]
    per_call = 2 if model == "gan" else 1
    all_steps = []
    for out in w_out:
        gs = _global_steps(out)
        assert gs, out
        assert gs == sorted(gs) and len(set(gs)) == len(gs)            # monotone per worker
        assert "Total Time:" in out
        all_steps += gs
    # shared counter: both workers advanced the same global_step; nobody repeats a value
    assert len(set(all_steps)) == len(all_steps)
    assert max(all_steps) >= steps and max(all_steps) < steps + 2 * per_call + 1
    if model == "gan":
        # both train ops carry global_step (gan:158-159): every call advances it by two
        assert all(b - a >= 2 for o in w_out for a, b in zip(_global_steps(o), _global_steps(o)[1:]))
    if model == "lstm":
        assert all(re.search(r"Test-Accuracy: \d\.\d{4}", o) for o in w_out)
    # chief checkpoints into model_dir with the reference's file naming
    files = os.listdir(model_dir)
    assert "checkpoint" in files and "graph.json" in files
    assert any(re.fullmatch(r"model\.ckpt-\d+\.index", f) for f in files)


def test_sync_mode_one_step_per_aggregate(tmp_path):
    ps_out, w_out, _ = _run_cluster("lstm", tmp_path, ["--num_steps=12", "--sync", "--save_model_secs=0"])
    a, b = _global_steps(w_out[0]), _global_steps(w_out[1])
    # every aggregate of 2 gradients advances global_step by one and both workers see the same token
    n = min(len(a), len(b))
    assert n >= 10 and a[:n] == b[:n] == list(range(1, n + 1))


def test_sync_mode_with_a_backup_worker_shuts_down_cleanly(tmp_path):
    """--replicas_to_aggregate=1 with two workers: aggregates of one gradient, the other worker's is
    dropped when it is stale.  When the chief reaches num_steps the worker still blocked on a token
    must be released (sentinel token), send its done token, and the ps must quit."""
    ps_out, w_out, _ = _run_cluster("lstm", tmp_path, ["--num_steps=15", "--sync", "--replicas_to_aggregate=1",
                                                       "--save_model_secs=0"])
    assert "ps 0: quitting" in ps_out[0] and "ps 0 received done 1" in ps_out[0]
    steps = _global_steps(w_out[0]) + _global_steps(w_out[1])
    assert max(steps) >= 15 and "Total Time" in w_out[0] and "Total Time" in w_out[1]


def test_two_ps_placement_and_resume(tmp_path):
    # first run writes checkpoints; second run (chief restarts) must resume from the latest one
    _, w1, model_dir = _run_cluster("lstm", tmp_path, ["--num_steps=20", "--save_model_secs=1"], n_ps=2)
    first_max = max(_global_steps(w1[0]) + _global_steps(w1[1]))
    from psb200.checkpoint import Saver, latest_checkpoint
    latest = latest_checkpoint(model_dir)
    assert latest is not None
    saved_step = int(Saver.restore(latest)["Variable_2"])
    assert saved_step >= 20 and set(Saver.list_variables(latest)) == {
        "Variable", "Variable_1", "rnn/basic_lstm_cell/kernel", "rnn/basic_lstm_cell/bias", "Variable_2"}
    _, w2, _ = _run_cluster("lstm", tmp_path, ["--num_steps=%d" % (saved_step + 10), "--save_model_secs=1"], n_ps=2)
    second = _global_steps(w2[0]) + _global_steps(w2[1])
    assert min(second) > saved_step - 1 and max(second) >= saved_step + 10     # continued, not restarted
    assert first_max >= 20


def test_launcher_runs_a_whole_cluster(tmp_path):
    from psb200.launch import launch
    rc, logs = launch("encoder", workers=2, ps=1, backend="cpu",
                      extra=["--synthetic", "--num_steps=10", "--save_model_secs=0"],
                      model_dir=str(tmp_path / "m"), quiet=True, timeout=120)
    assert rc == 0, logs
    assert "ps 0: quitting" in logs["ps:0"] and "Total Time:" in logs["worker:1"]


def test_tf_bundle_checkpoints_resume(tmp_path):
    """--checkpoint_format=tf: the chief writes TensorFlow V2 bundles and a restarted chief resumes from one."""
    from psb200 import tf_bundle
    from psb200.checkpoint import latest_checkpoint
    from psb200.launch import launch
    model_dir = str(tmp_path / "m")
    extra = ["--synthetic", "--save_model_secs=1", "--checkpoint_format=tf"]
    rc, logs = launch("lstm", workers=2, ps=1, backend="cpu", extra=extra + ["--num_steps=8"],
                      model_dir=model_dir, quiet=True, timeout=120)
    assert rc == 0, logs
    latest = latest_checkpoint(model_dir)
    assert latest is not None and tf_bundle.is_tf_bundle(latest)
    saved = int(tf_bundle.read_bundle(latest)["Variable_2"])
    assert saved >= 8
    rc, logs = launch("lstm", workers=2, ps=1, backend="cpu", extra=extra + ["--num_steps=%d" % (saved + 4)],
                      model_dir=model_dir, quiet=True, timeout=120)
    assert rc == 0, logs
    steps = _global_steps(logs["worker:0"]) + _global_steps(logs["worker:1"])
    assert min(steps) > saved - 1 and max(steps) >= saved + 4


def test_killed_worker_rejoins_and_cluster_finishes(tmp_path):
    """Fault injection: a non-chief worker is SIGKILLed mid-run and started again.  Async
    workers never wait for each other (device filters), so the chief keeps training; the restarted
    worker finds the variables initialised, continues from the shared global_step and sends the done
    token its first incarnation never sent, which lets the ps exit."""
    ports = _ports(3)
    ps = ["127.0.0.1:%d" % ports[0]]
    workers = ["127.0.0.1:%d" % p for p in ports[1:]]
    model_dir = str(tmp_path / "ckpt")
    extra = ["--num_steps=400", "--save_model_secs=0"]
    p_ps = _launch("encoder", "ps", 0, ps, workers, model_dir, extra)
    p_w0 = _launch("encoder", "worker", 0, ps, workers, model_dir, extra)
    p_w1 = _launch("encoder", "worker", 1, ps, workers, model_dir, extra)
    # let worker 1 make some progress, then kill it without any clean-up
    seen = ""
    deadline = time.time() + 120
    while "Local step 3 " not in seen and time.time() < deadline:
        line = p_w1.stdout.readline()
        if not line:
            break
        seen += line
    assert "Local step 3 " in seen, seen
    p_w1.kill()
    p_w1.wait()
    first_steps = _global_steps(seen)
    p_w1b = _launch("encoder", "worker", 1, ps, workers, model_dir, extra)
    outs = {}
    for name, p in (("w0", p_w0), ("w1b", p_w1b), ("ps", p_ps)):
        try:
            outs[name], _ = p.communicate(timeout=170)
        except subprocess.TimeoutExpired:
            for q in (p_ps, p_w0, p_w1b):
                q.kill()
            raise AssertionError("cluster did not finish after the restart: " + name)
        assert p.returncode == 0, outs[name][-2000:]
    again = _global_steps(outs["w1b"])
    assert again and min(again) > max(first_steps)            # continued from the shared counter, no re-init
    assert max(_global_steps(outs["w0"]) + again) >= 400
    assert "ps 0 received done 1" in outs["ps"] and "ps 0: quitting" in outs["ps"]
