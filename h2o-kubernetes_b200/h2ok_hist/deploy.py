"""SURVEY.md §8(f) rows 2-3, as far as they can be built here: the manifests `h2ok deploy` sends to the API server,
restated as plain data so that (i) `deploy --cluster_size N` without `--gpus` is structurally identical to what the
reference renders (src/k8s/templates.rs:9-57 StatefulSet, :74-89 Service, placeholder substitution :59-68, :91-97) and
(ii) `--gpus G` adds `nvidia.com/gpu` limits/requests and a node selector so the pods land on a B200 node.

No Kubernetes client, no cluster: this module only renders.  The reference's toolchain (cargo) is absent, so the
corresponding Rust edit (a `gpus` field with #[serde(default)] on both specification structs, src/cli/mod.rs:88-109 and
src/k8s/mod.rs:58-80, a `--gpus` clap arg next to `--cpus`, src/cli/mod.rs:207-212, and two conditional blocks in
src/k8s/templates.rs) is delivered as unverified source under reference_patch/; this module is the checked restatement
of what that patch renders.  `h2ok bench` is the C++ binary tools/h2ok-bench (same flags); the `bench` sub-command
here hands its arguments to it.
"""
import argparse
import os
import re

IMAGE, TAG = "h2oai/h2o-open-source-k8s", "latest"  # src/k8s/mod.rs:115-116
_QUANTITY = re.compile(r"^([+-]?[0-9.]+)([eEinumkKMGTP]*[-+]?[0-9]*)$")  # the reference's memory validator, cli/mod.rs:268-279


def stateful_set(name, namespace, nodes, memory_percentage=50, memory="1Gi", cpus=1, gpus=None,
                 gpu_node_selector=None):
    """templates.rs:9-57 with its quirks kept: serviceName is the literal `h2o-service` (templates.rs:16), cpu/memory
    request == limit, a single readiness probe on the leader-node endpoint."""
    resources = {"cpu": str(cpus), "memory": memory}
    limits, requests = dict(resources), dict(resources)
    pod_spec = {}
    if gpus:
        limits["nvidia.com/gpu"] = str(gpus)
        requests["nvidia.com/gpu"] = str(gpus)
        pod_spec["nodeSelector"] = dict(gpu_node_selector or {"nvidia.com/gpu.product": "NVIDIA-B200"})
    container = {
        "name": name,
        "image": f"{IMAGE}:{TAG}",
        "command": ["/bin/bash", "-c",
                    f"java -XX:+UseContainerSupport -XX:MaxRAMPercentage={memory_percentage} -jar /opt/h2oai/h2o-3/h2o.jar"],
        "ports": [{"containerPort": 54321, "protocol": "TCP"}],
        "readinessProbe": {"httpGet": {"path": "/kubernetes/isLeaderNode", "port": 8081}, "initialDelaySeconds": 5,
                           "periodSeconds": 5, "failureThreshold": 1},
        "resources": {"limits": limits, "requests": requests},
        "env": [
            {"name": "H2O_KUBERNETES_SERVICE_DNS", "value": f"{name}-service.{namespace}.svc.cluster.local"},
            {"name": "H2O_NODE_LOOKUP_TIMEOUT", "value": "180"},
            {"name": "H2O_NODE_EXPECTED_COUNT", "value": str(nodes)},
            {"name": "H2O_KUBERNETES_API_PORT", "value": "8081"},
        ],
    }
    pod_spec["containers"] = [container]
    return {
        "apiVersion": "apps/v1",
        "kind": "StatefulSet",
        "metadata": {"name": f"{name}-stateful-set", "namespace": namespace},
        "spec": {
            "serviceName": "h2o-service",
            "podManagementPolicy": "Parallel",
            "replicas": nodes,
            "selector": {"matchLabels": {"app": name}},
            "template": {"metadata": {"labels": {"app": name}}, "spec": pod_spec},
        },
    }


def service(name, namespace):
    """templates.rs:74-89: headless service in front of the pods."""
    return {
        "apiVersion": "v1",
        "kind": "Service",
        "metadata": {"name": f"{name}-service", "namespace": namespace},
        "spec": {"type": "ClusterIP", "clusterIP": "None", "selector": {"app": name},
                 "ports": [{"protocol": "TCP", "port": 80, "targetPort": 54321}]},
    }


def _positive_int(s):  # cli/mod.rs:245-252
    v = int(s)
    if v <= 0:
        raise argparse.ArgumentTypeError("must be greater than zero")
    return v


def _not_negative_int(s):
    v = int(s)
    if v < 0:
        raise argparse.ArgumentTypeError("must not be negative")
    return v


def _percentage(s):  # cli/mod.rs:259-266
    v = int(s)
    if not 1 <= v <= 100:
        raise argparse.ArgumentTypeError("must be between 1 and 100")
    return v


def _memory(s):  # cli/mod.rs:268-279
    if not _QUANTITY.match(s):
        raise argparse.ArgumentTypeError("not a Kubernetes quantity")
    return s


def build_parser():
    """The `deploy` flags of the reference (cli/mod.rs:166-213: snake_case long names, same defaults) plus `--gpus`,
    and a `bench` subcommand that forwards to bench.py."""
    ap = argparse.ArgumentParser(prog="h2ok")
    sub = ap.add_subparsers(dest="command", required=True)
    d = sub.add_parser("deploy")
    d.add_argument("--cluster_size", "-s", type=_positive_int, required=True)
    d.add_argument("--namespace", "-n", default="default")
    d.add_argument("--cluster_name", "-c", default="h2o-cluster")
    d.add_argument("--memory_percentage", "-p", type=_percentage, default=50)
    d.add_argument("--memory", "-m", type=_memory, default="1Gi")
    d.add_argument("--cpus", type=_positive_int, default=1)
    d.add_argument("--gpus", type=_positive_int, default=None, help="nvidia.com/gpu per pod (new)")
    b = sub.add_parser("bench")  # the flags of tools/h2ok_bench.cpp (and of reference_patch/src/cli/bench.rs)
    b.add_argument("--shape", choices=["c2", "c3", "c4"], default="c3")
    b.add_argument("--rows", "-r", type=_not_negative_int, default=0)
    b.add_argument("--cols", type=_not_negative_int, default=0)
    b.add_argument("--bins", "-b", type=_not_negative_int, default=0)
    b.add_argument("--depth", "-d", type=_not_negative_int, default=0)
    b.add_argument("--gpus", "-g", type=_positive_int, default=1)
    b.add_argument("--trees", "-t", type=_positive_int, default=3)
    b.add_argument("--seed", type=_not_negative_int, default=20261106)
    return ap


BENCH_BINARY = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "h2ok-bench")


def bench_argv(args):
    """The command line `h2ok bench ...` stands for: the stand-alone binary with the same flags."""
    return [BENCH_BINARY, "--shape", args.shape, "--rows", str(args.rows), "--cols", str(args.cols), "--bins", str(args.bins),
            "--depth", str(args.depth), "--gpus", str(args.gpus), "--trees", str(args.trees), "--seed", str(args.seed)]


def render(args):
    return [service(args.cluster_name, args.namespace),
            stateful_set(args.cluster_name, args.namespace, args.cluster_size, args.memory_percentage, args.memory,
                         args.cpus, args.gpus)]


def main(argv=None):
    """`deploy` prints the manifests it would send (YAML documents); `bench` runs tools/h2ok-bench."""
    args = build_parser().parse_args(argv)
    if args.command == "bench":
        if not os.path.exists(BENCH_BINARY):
            raise SystemExit(f"{BENCH_BINARY} is missing: build it with `make -C h2o-kubernetes_b200/csrc`")
        os.execv(BENCH_BINARY, bench_argv(args))
    import yaml

    print(yaml.safe_dump_all(render(args), sort_keys=False), end="")


if __name__ == "__main__":
    main()
