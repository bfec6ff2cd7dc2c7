#!/usr/bin/env bash
# End-to-end check of the Kubernetes launch path on a local kind cluster (CPU / gloo, 2 pods).
# Asserts the same artefacts the reference's e2e script does: both pods exit 0, rank 0 printed the
# run summary, and the host-mounted run directory holds checkpoints, logs, config and metadata,
# plus a non-empty MLflow SQLite file.
set -euo pipefail
cd "$(dirname "$0")/.."

CLUSTER="${KIND_CLUSTER:-llmtrain-e2e}"
IMAGE="${IMAGE:-llmtrain-b200:latest}"
TIMEOUT="${TIMEOUT:-300s}"
KEEP="${KEEP_CLUSTER:-0}"
fail() { echo "E2E FAIL: $*" >&2; exit 1; }
for tool in docker kind kubectl; do command -v "$tool" >/dev/null || fail "$tool not found"; done

cleanup() { [[ "$KEEP" == "1" ]] || kind delete cluster --name "$CLUSTER" >/dev/null 2>&1 || true; }
trap cleanup EXIT

rm -rf runs mlflow-k8s && mkdir -p runs mlflow-k8s
kind create cluster --name "$CLUSTER" --config k8s/kind-config.yaml
docker build -t "$IMAGE" -f k8s/Dockerfile --build-arg BASE="${BASE_IMAGE:-pytorch/pytorch:latest}" .
kind load docker-image "$IMAGE" --name "$CLUSTER"

kubectl apply -f k8s/rbac.yaml -f k8s/storage.yaml -f k8s/configmap-cpu.yaml -f k8s/service.yaml
sed -e 's/completions: 8/completions: 2/' -e 's/parallelism: 8/parallelism: 2/' \
    -e 's/value: "8"/value: "2"/' -e '/nvidia.com\/gpu/d' -e '/resources:/d' k8s/job.yaml | kubectl apply -f -
kubectl wait --for=condition=complete job/llmtrain --timeout="$TIMEOUT" \
  || { kubectl logs -l app=llmtrain --prefix --tail=50 || true; fail "job did not complete"; }

rank0="$(kubectl get pods -l batch.kubernetes.io/job-completion-index=0,job-name=llmtrain -o name | head -1)"
kubectl logs "$rank0" | grep -q "final_step=" || fail "rank 0 did not print a training summary"
for pod in $(kubectl get pods -l job-name=llmtrain -o name); do
  code="$(kubectl get "$pod" -o jsonpath='{.status.containerStatuses[0].state.terminated.exitCode}')"
  [[ "$code" == "0" ]] || fail "$pod exited with $code"
done

run_dir="$(find runs -mindepth 1 -maxdepth 1 -type d | head -1)"
[[ -n "$run_dir" ]] || fail "no run directory on the host"
[[ "$(find runs -mindepth 1 -maxdepth 1 -type d | wc -l)" == "1" ]] || fail "more than one run directory"
ls "$run_dir"/checkpoints/step_*.pt >/dev/null 2>&1 || fail "no checkpoints"
for f in logs/train.log config.yaml meta.json; do [[ -s "$run_dir/$f" ]] || fail "missing $f"; done
[[ -s mlflow-k8s/mlflow.db ]] || fail "MLflow database missing or empty"
echo "E2E OK: $run_dir"
