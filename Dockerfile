# p2pfl_b200: build + test image (B200 / sm_100a).  The kernels need CUDA >= 12.8 to target
# compute_100a; nothing is compiled for other architectures.
FROM nvidia/cuda:12.9.1-devel-ubuntu24.04

RUN apt-get update && apt-get install -y --no-install-recommends python3 python3-pip python3-venv git \
    && rm -rf /var/lib/apt/lists/*
RUN python3 -m venv /opt/venv
ENV PATH=/opt/venv/bin:$PATH CUDA_HOME=/usr/local/cuda

# PyTorch built against CUDA 12.8+ (Blackwell support), the control plane's few dependencies
RUN pip install --no-cache-dir torch --index-url https://download.pytorch.org/whl/cu128 \
    && pip install --no-cache-dir numpy grpcio protobuf psutil requests pytest

WORKDIR /opt/p2pfl_b200
COPY . .
# in-tree kernel library (nvcc cross-compiles without a GPU) + the package itself
RUN python build.py && pip install --no-cache-dir --no-deps -e .

# CPU tests at build time; the GPU tier needs a B200: `docker run --gpus all ... pytest -m gpu`
RUN python -m pytest tests -q -m "not gpu"
CMD ["python", "-m", "p2pfl_b200", "experiment", "list"]
