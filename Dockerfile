# Image for the host:port pipeline (counterpart of the reference's Dockerfile + docker-compose.yml profiles): one
# container per stage worker plus a coordinator, CPU stages by default.  Start from any image that has PyTorch; with an
# NVIDIA base image and `--gpus` the workers take `--gpu`.
ARG BASE=pytorch/pytorch:latest
FROM ${BASE}
WORKDIR /opt/dcnn_b200
COPY . .
# native host runtime (loaders, augmentation, codecs, CPU kernels, TCP transport); the CUDA library is built only when
# nvcc is present in the base image
RUN python -c "from dcnn_b200.build import build_host; print(build_host())" && \
    (command -v nvcc >/dev/null && python -c "from dcnn_b200.build import build; print(build())" || true)
ENV PYTHONPATH=/opt/dcnn_b200 PYTHONUNBUFFERED=1
CMD ["python", "examples/network_worker.py", "8001"]
