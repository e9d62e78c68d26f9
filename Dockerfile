# B200 image: CUDA 12.9 toolchain (nvcc targets sm_100a), torch >= 2.11 + cu128.
# The reference pins a CUDA 12.6 runtime image, which cannot target Blackwell.
ARG BASE=nvcr.io/nvidia/pytorch:25.06-py3
FROM ${BASE}

ENV DEBIAN_FRONTEND=noninteractive \
    PYTHONUNBUFFERED=1 \
    PYTHONPATH=/workspace

WORKDIR /workspace

COPY b200ddp ./b200ddp
COPY csrc ./csrc
COPY include ./include
COPY oracle ./oracle
COPY launch/entrypoint.sh train.py __graft_entry__.py ./

# Compile the sm_100a kernels in-tree (b200ddp/_native/libb200ddp.so).
RUN chmod +x entrypoint.sh && python -m b200ddp._native.build

# Topology defaults, same names as the reference image; a single 8xB200 box is
# REPLICAS=1 NPROC_PER_NODE=8 (no discovery service needed).
ENV REPLICAS=2 \
    MASTER_PORT=29500 \
    NPROC_PER_NODE=1

ENTRYPOINT ["./entrypoint.sh"]
