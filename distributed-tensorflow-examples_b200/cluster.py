"""ClusterSpec, Server and role helpers.

``ClusterSpec({"ps": [...], "worker": [...]})`` and ``Server(cluster, job_name, task_index)``
keep the call shape of ``tf.train.ClusterSpec`` / ``tf.train.Server`` (``encoder:89-98``).
A :class:`Server` starts the native runtime (``csrc/ps_server.cpp``) on this task's
``host:port``; in a ps process that runtime *is* the job, in a worker process it only serves
control traffic (ping/stats), mirroring the reference where every task starts a server.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Mapping, Optional, Sequence, Union

from .control.client import PsClient, ps_lib, split_hostport


class ClusterSpec:
    def __init__(self, jobs: Mapping[str, Union[Sequence[str], Mapping[int, str]]]):
        self._jobs: Dict[str, List[str]] = {}
        for job, tasks in jobs.items():
            if isinstance(tasks, Mapping):
                tasks = [tasks[i] for i in sorted(tasks)]
            self._jobs[job] = [str(t) for t in tasks]

    @property
    def jobs(self) -> List[str]:
        return sorted(self._jobs)

    def num_tasks(self, job: str) -> int:
        return len(self._jobs[job])

    def job_tasks(self, job: str) -> List[str]:
        return list(self._jobs[job])

    def task_address(self, job: str, index: int) -> str:
        return self._jobs[job][index]

    def as_dict(self) -> Dict[str, List[str]]:
        return {k: list(v) for k, v in self._jobs.items()}

    def __eq__(self, other) -> bool:
        return isinstance(other, ClusterSpec) and self._jobs == other._jobs

    def __repr__(self) -> str:
        return "ClusterSpec(%r)" % self._jobs


class DoneQueue:
    """``tf.FIFOQueue(workers, tf.int32, shared_name="done_queue<i>")`` on ``/job:ps/task:i``
    (``create_done_queue``, ``encoder:54-59``)."""

    def __init__(self, client: PsClient, index: int, capacity: int):
        self.client, self.name = client, "done_queue%d" % index
        client.queue_create(self.name, capacity)

    def enqueue(self, value: int = 1) -> None:
        self.client.queue_enqueue(self.name, value)

    def dequeue(self, timeout: float = 0.0) -> int:
        return self.client.queue_dequeue(self.name, timeout)


class Server:
    def __init__(self, cluster: ClusterSpec, job_name: str, task_index: int = 0, start: bool = True):
        if job_name not in cluster.jobs:
            raise ValueError("job_name must be one of %s, got %r" % (cluster.jobs, job_name))
        self.cluster, self.job_name, self.task_index = cluster, job_name, int(task_index)
        self.address = cluster.task_address(job_name, self.task_index)
        self._handle: Optional[int] = None
        self.port = split_hostport(self.address)[1]
        if start:
            self.start()

    @property
    def target(self) -> str:
        """Address a local session connects to (``server.target``, ``encoder:101,157``)."""
        return "127.0.0.1:%d" % self.port

    def start(self) -> None:
        if self._handle is not None:
            return
        host, port = split_hostport(self.address)
        bound = ctypes.c_int(0)
        h = ps_lib().psb_ps_start(b"0.0.0.0", port, ctypes.byref(bound))
        if h < 0:
            raise OSError("cannot start the parameter-server runtime on %s (code %d)" % (self.address, h))
        self._handle, self.port = h, bound.value

    def stop(self) -> None:
        if self._handle is not None:
            ps_lib().psb_ps_stop(self._handle)
            self._handle = None

    def join(self) -> None:
        """Block forever (``server.join()``; unused by the reference, kept for API parity)."""
        import time
        while self._handle is not None:
            time.sleep(1.0)

    def done_queue(self, index: Optional[int] = None, capacity: int = 1) -> DoneQueue:
        """This ps task's own done-queue, through a session on ``server.target`` (``encoder:101-102``)."""
        return DoneQueue(PsClient(self.target), self.task_index if index is None else index, capacity)


def device_filters(task_index: int) -> List[str]:
    """``ConfigProto(device_filters=['/job:ps', '/job:worker/task:<i>'])`` (``encoder:156``): the
    set of tasks a worker session may talk to — every ps, and itself."""
    return ["/job:ps", "/job:worker/task:%d" % task_index]
