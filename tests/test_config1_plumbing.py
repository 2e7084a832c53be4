"""BASELINE config 1 as a parity case: `h2ok deploy --cluster_size 3` manifest render + the CPU histogram oracle
(plumbing, no GPU).  The expected manifests are written out by hand from SURVEY.md §8(c) / src/k8s/templates.rs; the
reference has no render-only mode, so this is a structural golden, not a run of the reference binary."""
import copy

import numpy as np
import yaml

from h2ok_hist import abi, deploy


def expected(name="h2o-cluster", ns="default", nodes=3):
    res = {"cpu": "1", "memory": "1Gi"}
    sts = {
        "apiVersion": "apps/v1", "kind": "StatefulSet",
        "metadata": {"name": f"{name}-stateful-set", "namespace": ns},
        "spec": {
            "serviceName": "h2o-service", "podManagementPolicy": "Parallel", "replicas": nodes,
            "selector": {"matchLabels": {"app": name}},
            "template": {
                "metadata": {"labels": {"app": name}},
                "spec": {"containers": [{
                    "name": name, "image": "h2oai/h2o-open-source-k8s:latest",
                    "command": ["/bin/bash", "-c", "java -XX:+UseContainerSupport -XX:MaxRAMPercentage=50 -jar /opt/h2oai/h2o-3/h2o.jar"],
                    "ports": [{"containerPort": 54321, "protocol": "TCP"}],
                    "readinessProbe": {"httpGet": {"path": "/kubernetes/isLeaderNode", "port": 8081},
                                       "initialDelaySeconds": 5, "periodSeconds": 5, "failureThreshold": 1},
                    "resources": {"limits": dict(res), "requests": dict(res)},
                    "env": [{"name": "H2O_KUBERNETES_SERVICE_DNS", "value": f"{name}-service.{ns}.svc.cluster.local"},
                            {"name": "H2O_NODE_LOOKUP_TIMEOUT", "value": "180"},
                            {"name": "H2O_NODE_EXPECTED_COUNT", "value": str(nodes)},
                            {"name": "H2O_KUBERNETES_API_PORT", "value": "8081"}],
                }]},
            },
        },
    }
    svc = {"apiVersion": "v1", "kind": "Service", "metadata": {"name": f"{name}-service", "namespace": ns},
           "spec": {"type": "ClusterIP", "clusterIP": "None", "selector": {"app": name},
                    "ports": [{"protocol": "TCP", "port": 80, "targetPort": 54321}]}}
    return svc, sts


def test_deploy_cluster_size_3_renders_as_the_reference_does():
    args = deploy.build_parser().parse_args(["deploy", "--cluster_size", "3"])
    svc, sts = deploy.render(args)
    want_svc, want_sts = expected()
    assert svc == want_svc and sts == want_sts
    # round-trips through YAML (what the reference parses with serde_yaml) unchanged
    assert yaml.safe_load(yaml.safe_dump(sts)) == want_sts


def test_gpus_only_adds_gpu_scheduling():
    p = deploy.build_parser()
    base = deploy.render(p.parse_args(["deploy", "-s", "3", "-c", "h2o-a", "-n", "ml"]))[1]
    gpu = deploy.render(p.parse_args(["deploy", "-s", "3", "-c", "h2o-a", "-n", "ml", "--gpus", "8"]))[1]
    stripped = copy.deepcopy(gpu)
    c = stripped["spec"]["template"]["spec"]["containers"][0]["resources"]
    assert c["limits"].pop("nvidia.com/gpu") == "8" and c["requests"].pop("nvidia.com/gpu") == "8"
    assert stripped["spec"]["template"]["spec"].pop("nodeSelector") == {"nvidia.com/gpu.product": "NVIDIA-B200"}
    assert stripped == base  # nothing else changes


def test_validators_follow_the_reference():
    import pytest

    p = deploy.build_parser()
    for bad in (["deploy", "-s", "0"], ["deploy", "-s", "3", "-p", "101"], ["deploy", "-s", "3", "-m", "lots"],
                ["deploy"], ["deploy", "-s", "3", "--gpus", "0"]):
        with pytest.raises(SystemExit):
            p.parse_args(bad)
    assert p.parse_args(["deploy", "-s", "2", "-m", "256Mi", "--cpus", "2"]).memory == "256Mi"


def test_cpu_histogram_oracle_runs_the_plumbing_case(oracle):
    # the "CPU histogram oracle" half of config 1: one small tree end to end on the CPU
    rows, cols = 5000, 28
    X, y, _ = oracle.synth(abi.SHAPE_HIGGS28, 20261106, 0, rows, cols)
    levels, nid, leaf = oracle.train_tree(X, rows, y, None, abi.Params.make(nbins=256, max_depth=3))
    assert len(levels) == 3 and (leaf >= 0).all()
    assert levels[0]["hist"][0, 0, :, 0].sum() == rows
    assert np.isfinite(levels[0]["splits"][0]["se_left"])
