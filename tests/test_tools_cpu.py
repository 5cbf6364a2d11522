"""Deployment helpers: Kubernetes manifests carry the reference command line (README.md:2, */Dockerfile)."""
import io
import contextlib

import pytest
import yaml

from psb200.flags import parse_flags
from psb200.tools import k8s_manifest as km


def test_cpu_cluster_manifest_is_one_pod_and_service_per_task():
    docs = km.manifests("lstm", workers=3, ps=2, image="reg/psb200:1", extra=["--sync"], model_claim="ckpt")
    kinds = [d["kind"] for d in docs]
    assert kinds == ["Service"] * 5 + ["Pod"] * 5
    names = [d["metadata"]["name"] for d in docs if d["kind"] == "Pod"]
    assert names == ["lstm-ps-0", "lstm-ps-1", "lstm-worker-0", "lstm-worker-1", "lstm-worker-2"]
    for d in docs:
        if d["kind"] == "Service":       # headless, selects exactly its own task's pod
            assert d["spec"]["clusterIP"] == "None" and d["spec"]["selector"] == {"task": d["metadata"]["name"]}
    for d in docs:
        if d["kind"] != "Pod":
            continue
        cmd = d["spec"]["containers"][0]["command"]
        assert cmd[:2] == ["python", "examples/distributed_lstm.py"]
        fl = parse_flags(cmd[2:])        # the command line must parse with the entrypoint's own flag parser
        assert fl.ps_hosts == "lstm-ps-0:2222,lstm-ps-1:2222"
        assert fl.worker_hosts == "lstm-worker-0:2222,lstm-worker-1:2222,lstm-worker-2:2222"
        assert fl.workers == 3 and fl.ps == 2 and fl.sync and fl.synthetic and fl.extra == []
        job, idx = d["metadata"]["name"].split("-")[1:]
        assert fl.job_name == job and fl.task_index == int(idx)
        assert d["metadata"]["labels"]["task"] == d["metadata"]["name"]
        has_ckpt = any(v["name"] == "checkpoints" for v in d["spec"].get("volumes", []))
        assert has_ckpt == (d["metadata"]["name"] == "lstm-worker-0")      # only the chief writes checkpoints


def test_data_claim_mounts_mnist_instead_of_synthetic():
    docs = km.manifests("encoder", workers=1, ps=1, image="i", data_claim="mnist")
    pod = [d for d in docs if d["kind"] == "Pod"][-1]
    assert "--synthetic" not in pod["spec"]["containers"][0]["command"]
    assert pod["spec"]["containers"][0]["volumeMounts"][0] == {"name": "data", "mountPath": "/data_dir", "readOnly": True}


def test_nvl_manifest_is_one_gpu_pod_running_the_launcher():
    docs = km.manifests("gan", workers=8, ps=1, image="i", backend="nvl", extra=["--sync", "--num_steps=100"])
    assert len(docs) == 1 and docs[0]["kind"] == "Pod"
    c = docs[0]["spec"]["containers"][0]
    assert c["resources"]["limits"]["nvidia.com/gpu"] == 8
    assert c["command"][:3] == ["python", "-m", "psb200.launch"] and c["command"][-3:] == ["--sync", "--num_steps=100", "--synthetic"]
    from psb200.launch import main as _  # noqa: F401  (the module the pod runs must import)


def test_cli_prints_loadable_yaml():
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        assert km.main(["--model", "encoder", "--workers", "2", "--image", "img", "--", "--num_steps=5"]) == 0
    docs = list(yaml.safe_load_all(buf.getvalue()))
    assert len(docs) == 6 and docs[-1]["spec"]["containers"][0]["command"][-2:] == ["--num_steps=5", "--synthetic"]
    with pytest.raises(ValueError):
        km.manifests("lstm", workers=0, ps=1, image="i")
