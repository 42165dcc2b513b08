#!/bin/bash
# Cluster bring-up for b200dp: namespaces -> Loki/Promtail/Grafana -> MPI Operator -> MPIJob.
# Same four steps as the reference's deploy_stack.sh (deploy_stack.sh:1-103) with three differences:
#   * the Promtail JSON pipeline in deploy/promtail-pipeline.yaml is installed with the chart;
#   * the job is read from a manifest file (deploy/b200dp-mnist.yaml) instead of being inlined a second time;
#   * every step is idempotent and the script takes overrides from the environment.
set -euo pipefail

NAMESPACE=${NAMESPACE:-ml-ops}
LOKI_NAMESPACE=${LOKI_NAMESPACE:-loki}
MPI_OPERATOR_MANIFEST=${MPI_OPERATOR_MANIFEST:-https://raw.githubusercontent.com/kubeflow/mpi-operator/v0.6.0/deploy/v2beta1/mpi-operator.yaml}
JOB_MANIFEST=${JOB_MANIFEST:-$(dirname "$0")/b200dp-mnist.yaml}
LOKI_PVC_SIZE=${LOKI_PVC_SIZE:-5Gi}
HERE=$(cd "$(dirname "$0")" && pwd)

echo "[1/4] namespaces: $NAMESPACE $LOKI_NAMESPACE"
for ns in "$NAMESPACE" "$LOKI_NAMESPACE"; do
  kubectl get namespace "$ns" >/dev/null 2>&1 || kubectl create namespace "$ns"
done

echo "[2/4] Grafana Loki stack (Loki + Promtail + Grafana) with the b200dp JSON pipeline"
helm repo add grafana https://grafana.github.io/helm-charts >/dev/null
helm repo update >/dev/null
helm upgrade --install loki grafana/loki-stack \
  --namespace "$LOKI_NAMESPACE" \
  --set grafana.enabled=true \
  --set promtail.enabled=true \
  --set loki.persistence.enabled=true \
  --set loki.persistence.size="$LOKI_PVC_SIZE" \
  -f "$HERE/promtail-pipeline.yaml" \
  --wait

# Grafana's dashboard sidecar (enabled in the chart by default) loads every ConfigMap labelled grafana_dashboard=1
kubectl create configmap b200dp-dashboard -n "$LOKI_NAMESPACE" --from-file=b200dp.json="$HERE/grafana-dashboard.json" \
  --dry-run=client -o yaml | kubectl label --local -f - grafana_dashboard=1 -o yaml | kubectl apply -f -
# alert rules for the Loki ruler (stall, rank error, restart, straggler)
kubectl create configmap b200dp-loki-rules -n "$LOKI_NAMESPACE" --from-file=b200dp-rules.yaml="$HERE/loki-rules.yaml" \
  --dry-run=client -o yaml | kubectl apply -f -

echo "[3/4] kubeflow MPI Operator (v2beta1)"
kubectl apply --server-side -f "$MPI_OPERATOR_MANIFEST"

echo "[4/4] MPIJob from $JOB_MANIFEST"
kubectl apply -n "$NAMESPACE" -f "$JOB_MANIFEST"

echo "done. follow the launcher:  kubectl logs -n $NAMESPACE -f -l training.kubeflow.org/job-role=launcher"
