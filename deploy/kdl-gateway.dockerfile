# One-container variant: the /predict route (model_server.py:59-70) with the engine in process -- no gRPC hop, no 1 MB float32
# tensor on the wire, concurrent requests coalesced into one forward call.  Port 9696 like gateway.dockerfile.
# Build from the repository root:   docker build -f deploy/kdl-gateway.dockerfile -t kdl-gateway .
FROM kdl-serving
ENV KDL_ENABLE_BATCHING=1
ENV KDL_BATCH_TIMEOUT_MICROS=1000
EXPOSE 9696
ENTRYPOINT ["python3", "-m", "kdl_b200.http_gateway"]
