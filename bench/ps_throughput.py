"""Loopback throughput of the native TCP parameter server (csrc/ps_server.cpp): variable reads
(parameter pull), Hogwild applies (gradient push + optimizer on the ps) and queue round trips.

    python bench/ps_throughput.py [--mb 4] [--iters 200] [--clients 2]

CPU only.  Prints one JSON line.
"""
import argparse
import json
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from psb200.cluster import ClusterSpec, Server  # noqa: E402
from psb200.control.client import PsClient  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mb", type=float, default=1.9, help="variable size in MB (encoder model = 1.87 MB)")
ap.add_argument("--iters", type=int, default=200)
ap.add_argument("--clients", type=int, default=2)
a = ap.parse_args()

cluster = ClusterSpec({"ps": ["127.0.0.1:0"], "worker": ["127.0.0.1:0"]})
server = Server(cluster, job_name="ps", task_index=0)
server.start()
addr = server.target
n = int(a.mb * (1 << 20) / 4)
c0 = PsClient(addr)
c0.var_declare("w", [n], torch.float32)
c0.var_assign("w", torch.zeros(n))
c0.queue_create("q", 1)
results = {}


def worker(idx):
    c = PsClient(addr)
    g = torch.full((n,), 1e-3)
    t0 = time.perf_counter()
    for _ in range(a.iters):
        c.var_read("w", [n], torch.float32)
    t1 = time.perf_counter()
    for _ in range(a.iters):
        c.apply("w", g, 0, [0.01, 0.0, 0.0, 0.0], [])
    t2 = time.perf_counter()
    results[idx] = (t1 - t0, t2 - t1)
    c.close()


threads = [threading.Thread(target=worker, args=(i,)) for i in range(a.clients)]
for t in threads:
    t.start()
for t in threads:
    t.join()
t0 = time.perf_counter()
for _ in range(2000):
    c0.queue_enqueue("q", 1)
    c0.queue_dequeue("q")
rt = (time.perf_counter() - t0) / 4000
final = c0.var_read("w", [n], torch.float32)
expect = -0.01 * 1e-3 * a.iters * a.clients
read_s = max(r[0] for r in results.values())
apply_s = max(r[1] for r in results.values())
mb = n * 4 / 1e6
print(json.dumps({
    "variable_mb": round(mb, 2), "clients": a.clients, "iters": a.iters,
    "pull_ms": round(1e3 * read_s / a.iters, 3), "pull_MBps_total": round(mb * a.iters * a.clients / read_s, 1),
    "push_apply_ms": round(1e3 * apply_s / a.iters, 3), "push_MBps_total": round(mb * a.iters * a.clients / apply_s, 1),
    "queue_round_trip_us": round(1e6 * rt, 1),
    "hogwild_lost_update_fraction": round(1.0 - float(final.mean()) / expect, 4),
}))
c0.close()
server.stop()
