"""deploy/ (SURVEY.md §8f-4): the manifests keep the reference's names and ports (serving-gateway-deployment.yaml:23-24,
tf-serving-clothing-model-service.yaml:9-10, serving-gateway-service.yaml:10-11) and ask for a GPU.  CPU only."""
import os

import pytest

yaml = pytest.importorskip("yaml")
D = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "deploy")


def _load(name, kind):
    with open(os.path.join(D, name)) as f:
        docs = [d for d in yaml.safe_load_all(f) if d]
    (doc,) = [d for d in docs if d["kind"] == kind]
    return doc


def test_model_server_deployment_and_service_keep_the_reference_contract():
    dep = _load("kdl-model-server.yaml", "Deployment")
    svc = _load("kdl-model-server.yaml", "Service")
    assert dep["kind"] == "Deployment" and dep["metadata"]["name"] == "tf-serving-clothing-model"
    c = dep["spec"]["template"]["spec"]["containers"][0]
    assert c["ports"] == [{"containerPort": 8500}]
    assert c["resources"]["limits"]["nvidia.com/gpu"] == 1
    assert c["readinessProbe"]["tcpSocket"]["port"] == 8500
    assert dep["spec"]["selector"]["matchLabels"] == dep["spec"]["template"]["metadata"]["labels"] == svc["spec"]["selector"]
    # the unmodified gateway dials tf-serving-clothing-model.default.svc.cluster.local:8500
    assert svc["metadata"]["name"] == "tf-serving-clothing-model"
    assert svc["spec"]["ports"][0]["port"] == 8500 and svc["spec"]["ports"][0]["targetPort"] == 8500


def test_single_container_gateway_keeps_the_public_service():
    dep = _load("kdl-gateway-inprocess.yaml", "Deployment")
    svc = _load("kdl-gateway-inprocess.yaml", "Service")
    c = dep["spec"]["template"]["spec"]["containers"][0]
    assert c["ports"] == [{"containerPort": 9696}] and c["resources"]["limits"]["nvidia.com/gpu"] == 1
    assert c["readinessProbe"]["httpGet"] == {"path": "/healthz", "port": 9696}
    assert svc["metadata"]["name"] == "serving-gateway" and svc["spec"]["type"] == "LoadBalancer"
    assert svc["spec"]["ports"][0]["port"] == 80 and svc["spec"]["ports"][0]["targetPort"] == 9696
    assert svc["spec"]["selector"] == dep["spec"]["template"]["metadata"]["labels"]


def test_dockerfiles_start_the_module_entry_points():
    s = open(os.path.join(D, "kdl-serving.dockerfile")).read()
    assert 'ENTRYPOINT ["python3", "-m", "kdl_b200.serving"]' in s and "EXPOSE 8500" in s and "MODEL_NAME=clothing-model" in s
    g = open(os.path.join(D, "kdl-gateway.dockerfile")).read()
    assert 'ENTRYPOINT ["python3", "-m", "kdl_b200.http_gateway"]' in g and "EXPOSE 9696" in g
