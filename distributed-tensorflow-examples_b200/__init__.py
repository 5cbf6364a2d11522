"""psb200 — a B200-native parameter-server training framework.

Capabilities mirror the three TF1 scripts of Xingskcs/Distributed-TensorFlow-Examples
(``encoder/distributed_encoder.py``, ``gan/distributed_gan.py``,
``lstm/distributed_lstm.py``): ClusterSpec ps/worker roles, round-robin variable
placement, async (Hogwild) and sync (aggregate R-of-N) updates, chief-driven
init/checkpoint/resume, done-queue shutdown and the three model entrypoints.

Two transports carry the parameter traffic:

* ``store.tcp``  — a native C++ parameter server reached over loopback/ethernet TCP
  (the "1 ps + 2 workers on CPU localhost" plumbing configuration);
* ``store.nvl``  — one process per B200; every GPU owns a parameter shard in HBM and
  the gradient push / parameter pull run inside hand-written sm_100a kernels over
  NVLink peer memory.

Import name is ``psb200`` (see ``/psb200/__init__.py``).
"""

__version__ = "0.1.0"

import os as _os

PACKAGE_DIR = _os.path.dirname(_os.path.abspath(__file__))
REPO_ROOT = _os.path.dirname(PACKAGE_DIR)
