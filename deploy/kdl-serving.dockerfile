# GPU replacement for the reference's TF-Serving image (tf-serving.dockerfile: tensorflow/serving:2.3.0 + the SavedModel under
# /models/clothing-model/1).  Same port (8500), same RPC (tensorflow.serving.PredictionService/Predict), same model name.
# Build from the repository root:   docker build -f deploy/kdl-serving.dockerfile -t kdl-serving .
# The CUDA library is compiled for sm_100a at image build time (nvcc cross-compiles; no GPU needed to build).
FROM nvidia/cuda:12.9.1-devel-ubuntu24.04 AS build
RUN apt-get update && apt-get install -y --no-install-recommends make g++ && rm -rf /var/lib/apt/lists/*
WORKDIR /src
COPY include include
COPY kubernetes-deep-learning_b200/csrc kubernetes-deep-learning_b200/csrc
RUN make -C kubernetes-deep-learning_b200/csrc

FROM nvidia/cuda:12.9.1-runtime-ubuntu24.04
ENV PYTHONUNBUFFERED=TRUE
RUN apt-get update && apt-get install -y --no-install-recommends python3 python3-pip && rm -rf /var/lib/apt/lists/*
# torch is only plumbing here (device memory, streams); grpcio + protobuf carry the wire format
RUN pip3 install --no-cache-dir --break-system-packages torch numpy grpcio protobuf pillow
WORKDIR /app
COPY kdl_b200.py kdl_b200.py
COPY include include
COPY kubernetes-deep-learning_b200 kubernetes-deep-learning_b200
COPY --from=build /src/kubernetes-deep-learning_b200/csrc/libkdl.so kubernetes-deep-learning_b200/csrc/libkdl.so
# the trained weights: xception_v4_large_08_0.894.h5 of convert.py:4 (Keras HDF5, read by kdl_b200.keras_io) -- the counterpart of
# `COPY clothing-model /models/clothing-model/1`
ENV MODEL_NAME=clothing-model
ENV KDL_WEIGHTS=/models/clothing-model/1/xception_v4_large_08_0.894.h5
COPY xception_v4_large_08_0.894.h5 /models/clothing-model/1/xception_v4_large_08_0.894.h5
EXPOSE 8500
ENTRYPOINT ["python3", "-m", "kdl_b200.serving"]
