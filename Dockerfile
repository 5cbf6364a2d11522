# Container recipe equivalent to the reference's */Dockerfile (FROM tensorflow:1.11.0, COPY the
# script, COPY mnist_dataset /data_dir).  One image serves all three entrypoints; an external
# orchestrator (or `python -m psb200.launch`) runs one task per container/GPU.
FROM nvcr.io/nvidia/pytorch:25.01-py3
COPY . /opt/psb200
WORKDIR /opt/psb200
RUN python -m psb200.build
# COPY mnist_dataset /data_dir      # the four MNIST idx(.gz) files; or run with --synthetic
# e.g.  python examples/distributed_lstm.py --ps_hosts=... --worker_hosts=... --job_name=worker --task_index=0
