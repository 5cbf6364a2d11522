#!/usr/bin/env python
"""distributed_gan — same command line as the reference's ``gan/distributed_gan.py``.

    python examples/distributed_gan.py --ps_hosts=h:p[,..] --worker_hosts=h:p[,..] \\
        --job_name=ps|worker --task_index=i [--workers=N] [--data_dir=..] [--model_dir=..]

Extensions (all default to the reference behaviour): --sync, --replicas_to_aggregate,
--backend=cpu|nvl, --synthetic, --num_steps, --save_model_secs, --gan_update, --reinit_on_join.
Model hyper-parameters are module constants of ``psb200.models.gan``, as in the reference.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from psb200 import runner  # noqa: E402
from psb200.flags import parse_flags  # noqa: E402


def main(argv=None):
    FLAGS = parse_flags(argv, prog="distributed_gan")
    return runner.main("gan", FLAGS)


if __name__ == "__main__":
    main()
