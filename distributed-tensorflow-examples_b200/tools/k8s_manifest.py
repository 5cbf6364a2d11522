"""Kubernetes manifests for a ps/worker cluster of one of the examples.

The reference calls itself "Distributed TensorFlow Examples for Kubernetes" (``README.md:2``) and
ships one three-line Dockerfile per example (``encoder/Dockerfile:1-3``) but no manifests: every
task is ``python /distributed_X.py --ps_hosts=... --job_name=... --task_index=i`` started by hand
or by an orchestrator.  This tool writes what that orchestrator needs.

    python -m psb200.tools.k8s_manifest --model lstm --workers 3 --ps 1 --image registry/psb200:latest

``--backend=cpu`` (default): one Pod + one headless Service per task (stable DNS names feed
``--ps_hosts/--worker_hosts``), like the reference's one-process-per-task topology.
``--backend=nvl``: the workers share CUDA-IPC peer memory, so they live in ONE Pod that requests
``--workers`` GPUs and runs ``python -m psb200.launch`` (ps tasks + one worker process per GPU).
"""
from __future__ import annotations

import argparse
import sys
from typing import Dict, List, Optional

import yaml

MODELS = ("encoder", "gan", "lstm")


def _task_name(model: str, job: str, index: int) -> str:
    return "%s-%s-%d" % (model, job, index)


def _service(name: str, port: int, namespace: str, labels: Dict[str, str]) -> dict:
    return {"apiVersion": "v1", "kind": "Service",
            "metadata": {"name": name, "namespace": namespace, "labels": labels},
            "spec": {"clusterIP": "None", "selector": {"task": name},
                     "ports": [{"name": "psb200", "port": port, "targetPort": port}]}}


def _pod(name: str, namespace: str, labels: Dict[str, str], image: str, command: List[str], port: int,
         gpus: int = 0, data_claim: Optional[str] = None, model_claim: Optional[str] = None) -> dict:
    container = {"name": "task", "image": image, "command": command, "workingDir": "/opt/psb200",
                 "ports": [{"containerPort": port}]}
    if gpus:
        container["resources"] = {"limits": {"nvidia.com/gpu": gpus}}
    mounts, volumes = [], []
    if data_claim:
        mounts.append({"name": "data", "mountPath": "/data_dir", "readOnly": True})
        volumes.append({"name": "data", "persistentVolumeClaim": {"claimName": data_claim}})
    if model_claim:
        mounts.append({"name": "checkpoints", "mountPath": "/tmp/checkpoints"})
        volumes.append({"name": "checkpoints", "persistentVolumeClaim": {"claimName": model_claim}})
    if mounts:
        container["volumeMounts"] = mounts
    spec = {"restartPolicy": "OnFailure", "containers": [container]}
    if volumes:
        spec["volumes"] = volumes
    return {"apiVersion": "v1", "kind": "Pod",
            "metadata": {"name": name, "namespace": namespace, "labels": dict(labels, task=name)},
            "spec": spec}


def manifests(model: str, workers: int, ps: int, image: str, backend: str = "cpu", port: int = 2222,
              namespace: str = "default", extra: Optional[List[str]] = None, data_claim: Optional[str] = None,
              model_claim: Optional[str] = None) -> List[dict]:
    """The list of Kubernetes objects for one cluster (Services first, then Pods)."""
    if model not in MODELS:
        raise ValueError("model must be one of %s" % (MODELS,))
    if workers < 1 or ps < 1:
        raise ValueError("need at least one ps and one worker")
    extra = list(extra or [])
    if not data_claim and "--synthetic" not in extra:
        extra.append("--synthetic")          # no MNIST volume: the image has no /data_dir (encoder/Dockerfile:3)
    labels = {"app": "psb200", "model": model}
    ps_names = [_task_name(model, "ps", i) for i in range(ps)]
    ps_hosts = ",".join("%s:%d" % (n, port) for n in ps_names)
    out: List[dict] = []
    script = "examples/distributed_%s.py" % model
    if backend == "cpu":
        w_names = [_task_name(model, "worker", i) for i in range(workers)]
        w_hosts = ",".join("%s:%d" % (n, port) for n in w_names)
        out += [_service(n, port, namespace, labels) for n in ps_names + w_names]
        common = ["--ps_hosts=" + ps_hosts, "--worker_hosts=" + w_hosts, "--workers=%d" % workers, "--ps=%d" % ps]
        for job, names in (("ps", ps_names), ("worker", w_names)):
            for i, n in enumerate(names):
                cmd = ["python", script] + common + ["--job_name=" + job, "--task_index=%d" % i] + extra
                out.append(_pod(n, namespace, labels, image, cmd, port, data_claim=data_claim,
                                model_claim=model_claim if (job, i) == ("worker", 0) else None))
    elif backend == "nvl":
        # The workers exchange CUDA-IPC handles, so the whole cluster is one Pod with `workers` GPUs;
        # psb200.launch starts the ps tasks (done-queues only) and one worker per GPU inside it.
        n = _task_name(model, "cluster", 0)
        cmd = ["python", "-m", "psb200.launch", "--model", model, "--workers", str(workers), "--ps", str(ps),
               "--backend", "nvl", "--model_dir", "/tmp/checkpoints", "--"] + extra
        out = [_pod(n, namespace, labels, image, cmd, port, gpus=workers, data_claim=data_claim, model_claim=model_claim)]
    else:
        raise ValueError("backend must be 'cpu' or 'nvl'")
    return out


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(prog="k8s_manifest", description=__doc__.split("\n\n")[0])
    ap.add_argument("--model", required=True, choices=MODELS)
    ap.add_argument("--workers", type=int, default=3)
    ap.add_argument("--ps", type=int, default=1)
    ap.add_argument("--image", required=True)
    ap.add_argument("--backend", default="cpu", choices=("cpu", "nvl"))
    ap.add_argument("--port", type=int, default=2222)
    ap.add_argument("--namespace", default="default")
    ap.add_argument("--data_claim", default="", help="PersistentVolumeClaim with the MNIST idx files (mounted at /data_dir)")
    ap.add_argument("--model_claim", default="", help="PersistentVolumeClaim for checkpoints (chief, /tmp/checkpoints)")
    ap.add_argument("extra", nargs="*", help="flags passed through to the entrypoint after '--'")
    a = ap.parse_args(argv)
    docs = manifests(a.model, a.workers, a.ps, a.image, a.backend, a.port, a.namespace, a.extra,
                     a.data_claim or None, a.model_claim or None)
    yaml.safe_dump_all(docs, sys.stdout, sort_keys=False)
    return 0


if __name__ == "__main__":
    sys.exit(main())
