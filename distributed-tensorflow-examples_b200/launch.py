"""Localhost launcher — the packaging/deployment layer.

The reference ships three Dockerfiles (``encoder/Dockerfile:1-3`` etc.) and leaves the
orchestration ("python /distributed_X.py --flags" once per task) to Kubernetes.  On one
8xB200 box the equivalent is one command that starts every task of the cluster:

    python -m psb200.launch --model lstm --workers 8 --ps 1 --backend nvl -- --sync --synthetic --num_steps 1000

It picks free ports, builds ``--ps_hosts/--worker_hosts``, starts ``ps`` and ``worker`` processes
of ``examples/distributed_<model>.py`` with the reference flags (worker i pinned to GPU i for
``--backend nvl``), streams their output with a ``[job:index]`` prefix and returns the first
non-zero exit code.
"""
from __future__ import annotations

import argparse
import os
import socket
import subprocess
import sys
import threading
from typing import List

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def free_ports(n: int) -> List[int]:
    socks = [socket.socket() for _ in range(n)]
    for s in socks:
        s.bind(("127.0.0.1", 0))
    ports = [s.getsockname()[1] for s in socks]
    for s in socks:
        s.close()
    return ports


def _pump(prefix: str, proc: subprocess.Popen, sink: List[str]) -> None:
    for line in proc.stdout:
        sink.append(line)
        sys.stdout.write("[%s] %s" % (prefix, line))
        sys.stdout.flush()


def launch(model: str, workers: int, ps: int, backend: str, extra: List[str], model_dir: str,
           quiet: bool = False, timeout: float = 0.0):
    ports = free_ports(ps + workers + 1)
    ps_hosts = ",".join("127.0.0.1:%d" % p for p in ports[:ps])
    worker_hosts = ",".join("127.0.0.1:%d" % p for p in ports[ps:ps + workers])
    script = os.path.join(ROOT, "examples", "distributed_%s.py" % model)
    procs, threads, logs = [], [], {}
    for job, n in (("ps", ps), ("worker", workers)):
        for i in range(n):
            cmd = [sys.executable, script, "--ps_hosts=" + ps_hosts, "--worker_hosts=" + worker_hosts,
                   "--job_name=" + job, "--task_index=%d" % i, "--workers=%d" % workers, "--ps=%d" % ps,
                   "--model_dir=" + model_dir, "--backend=" + backend] + extra
            env = dict(os.environ, PSB200_RDZV_PORT=str(ports[-1]))
            if backend == "nvl" and job == "worker":
                env["LOCAL_RANK"] = str(i)
            # many tasks share one host: keep each task's CPU thread pool small (torchrun does the same)
            env.setdefault("OMP_NUM_THREADS", "2" if (job == "ps" or backend == "cpu") else "4")
            p = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=ROOT, env=env)
            key = "%s:%d" % (job, i)
            logs[key] = []
            t = threading.Thread(target=_pump if not quiet else (lambda k, pr, s: s.extend(pr.stdout)),
                                 args=(key, p, logs[key]), daemon=True)
            t.start()
            procs.append((key, p))
            threads.append(t)
    rc = 0
    for key, p in procs:
        try:
            p.wait(timeout=timeout or None)
        except subprocess.TimeoutExpired:
            for _, q in procs:
                q.kill()
            rc = rc or 124
        if p.returncode and not rc:
            rc = p.returncode
    for t in threads:
        t.join(timeout=5)
    return rc, {k: "".join(v) for k, v in logs.items()}


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="psb200.launch")
    ap.add_argument("--model", required=True, choices=["encoder", "gan", "lstm"])
    ap.add_argument("--workers", type=int, default=2)
    ap.add_argument("--ps", type=int, default=1)
    ap.add_argument("--backend", default="cpu", choices=["cpu", "nvl"])
    ap.add_argument("--model_dir", default="/tmp/checkpoints")
    ap.add_argument("--timeout", type=float, default=0.0)
    ap.add_argument("rest", nargs=argparse.REMAINDER, help="flags passed through to every task (after --)")
    a = ap.parse_args(argv)
    extra = [x for x in a.rest if x != "--"]
    rc, _ = launch(a.model, a.workers, a.ps, a.backend, extra, a.model_dir, timeout=a.timeout)
    return rc


if __name__ == "__main__":
    sys.exit(main())
