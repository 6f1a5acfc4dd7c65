# Container recipe (the reference's Dockerfile:1-10 patches Horovod 0.18.2 into a TF/torch image;
# nothing needs patching here: the Horovod-style API is part of the package).
FROM nvcr.io/nvidia/pytorch:25.03-py3
COPY . /workspace/grace_b200
WORKDIR /workspace/grace_b200
RUN bash grace_b200/csrc/build.sh && pip install --no-deps -e .
