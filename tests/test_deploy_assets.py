"""The cluster-facing assets stay usable (W19): manifests parse, and the launcher understands their argv."""
import os
import subprocess

import yaml

from b200dp.launch import b200run as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
D = os.path.join(ROOT, "deploy")


def _job(name):
    return yaml.safe_load(open(os.path.join(D, name)))


def test_gpu_mpijob_manifest():
    j = _job("b200dp-mnist.yaml")
    assert j["apiVersion"] == "kubeflow.org/v2beta1" and j["kind"] == "MPIJob"
    spec = j["spec"]
    assert spec["runPolicy"]["cleanPodPolicy"] == "Running"
    launcher = spec["mpiReplicaSpecs"]["Launcher"]["template"]["spec"]["containers"][0]
    worker = spec["mpiReplicaSpecs"]["Worker"]
    a = L.build_parser().parse_args(launcher["args"])
    assert launcher["command"] == ["b200run"]
    # the three places that must agree on the world size
    assert a.np == spec["slotsPerWorker"] * worker["replicas"] == 8
    assert worker["template"]["spec"]["containers"][0]["resources"]["limits"]["nvidia.com/gpu"] == spec["slotsPerWorker"]
    assert a.command[:2] == ["python", "/examples/tensorflow_mnist_gpu.py"] and a.max_restarts == 2


def test_cpu_manifest_keeps_the_reference_argv():
    j = _job("b200dp-mnist-cpu.yaml")
    launcher = j["spec"]["mpiReplicaSpecs"]["Launcher"]["template"]["spec"]["containers"][0]
    ref = ["-np", "2", "--allow-run-as-root", "-bind-to", "none", "-map-by", "slot", "-x", "LD_LIBRARY_PATH", "-x", "PATH",
           "-mca", "pml", "ob1", "-mca", "btl", "^openib"]                      # horovod/tensorflow-mnist.yaml:19-36
    assert launcher["command"] == ["mpirun"] and launcher["args"][:len(ref)] == ref
    a = L.build_parser().parse_args(launcher["args"])
    assert a.np == 2 and a.cpu and a.command[-2:] == ["/examples/tensorflow_mnist.py", "--synthetic"]
    assert j["spec"]["mpiReplicaSpecs"]["Worker"]["replicas"] * j["spec"]["slotsPerWorker"] == a.np


def test_promtail_pipeline_lifts_json_fields():
    p = yaml.safe_load(open(os.path.join(D, "promtail-pipeline.yaml")))
    stages = p["promtail"]["config"]["snippets"]["pipelineStages"]
    inner = [s for s in stages if "match" in s][0]["match"]["stages"]
    exprs = [s for s in inner if "json" in s][0]["json"]["expressions"]
    assert {"ts", "level", "event", "rank", "step", "loss"} <= set(exprs)


def test_deploy_script_and_dockerfile():
    assert subprocess.run(["bash", "-n", os.path.join(D, "deploy_stack.sh")]).returncode == 0
    sh = open(os.path.join(D, "deploy_stack.sh")).read()
    for needle in ("grafana/loki-stack", "promtail.enabled=true", "loki.persistence.size", "mpi-operator", "kubectl apply -n"):
        assert needle in sh
    df = open(os.path.join(D, "Dockerfile")).read()
    assert "tools/build.py" in df and "openssh-server" in df and "/examples/tensorflow_mnist.py" in df


def test_grafana_dashboard_and_loki_rules_use_events_the_framework_emits():
    """Every LogQL selector in the dashboard / alert rules names an event that some b200dp module really logs."""
    import json
    import re
    root = os.path.dirname(D)
    emitted = set()
    for dp, _, fns in os.walk(os.path.join(root, "b200dp")):
        for fn in fns:
            if fn.endswith(".py"):
                emitted |= set(re.findall(r'(?:info|warning|error|log)\("([a-z_]+)"', open(os.path.join(dp, fn)).read()))
    dash = json.load(open(os.path.join(D, "grafana-dashboard.json")))
    exprs = [t["expr"] for p in dash["panels"] for t in p["targets"]]
    assert len(dash["panels"]) >= 6 and dash["uid"] == "b200dp-train"
    rules = yaml.safe_load(open(os.path.join(D, "loki-rules.yaml")))
    alerts = rules["groups"][0]["rules"]
    assert {a["alert"] for a in alerts} >= {"B200dpTrainingStalled", "B200dpRankError", "B200dpJobRestarted", "B200dpStragglerRank"}
    exprs += [a["expr"] for a in alerts]
    used = set()
    for e in exprs:
        for m in re.findall(r'event=~?"([^"]+)"', e):
            used |= {x for x in m.split("|") if re.fullmatch(r"[a-z_]+", x)}
    assert used and used <= emitted, used - emitted
    sh = open(os.path.join(D, "deploy_stack.sh")).read()
    assert "grafana-dashboard.json" in sh and "grafana_dashboard=1" in sh and "loki-rules.yaml" in sh
    j = yaml.safe_load(open(os.path.join(D, "b200dp-mnist.yaml")))
    w = j["spec"]["mpiReplicaSpecs"]["Worker"]["template"]
    assert w["metadata"]["annotations"]["prometheus.io/port"] == "9400"
    assert "B200DP_METRICS_PORT=9400" in j["spec"]["mpiReplicaSpecs"]["Launcher"]["template"]["spec"]["containers"][0]["args"]
