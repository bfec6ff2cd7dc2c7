#!/usr/bin/env bash
# Rank bootstrap for a Kubernetes *Indexed* Job (one pod = one rank).
#
# Same launch protocol as the reference's entrypoint (k8s/entrypoint.sh:25-89 there): the pod's
# completion index is its global rank, rank 0 advertises its own pod IP as the rendezvous address,
# every other rank asks the API server for the IP of the index-0 pod of the same Job.  What is
# new for GPUs: LOCAL_RANK is NOT the global rank — it is the index of the GPU this rank should
# bind inside its own pod (0 when each pod requests one nvidia.com/gpu; rank % GPUS_PER_POD when a
# pod exposes several and runs one rank per GPU through torchrun instead of this script).
set -euo pipefail

: "${JOB_COMPLETION_INDEX:?this script expects an Indexed Job (JOB_COMPLETION_INDEX unset)}"
: "${WORLD_SIZE:?set WORLD_SIZE to the Job's completions}"
: "${JOB_NAME:?set JOB_NAME (used to find the rank-0 pod)}"
export MASTER_PORT="${MASTER_PORT:-29500}"
export RANK="${JOB_COMPLETION_INDEX}"
export LOCAL_RANK="${LOCAL_RANK:-$(( RANK % ${GPUS_PER_POD:-1} ))}"
CONFIG="${LLMTRAIN_CONFIG:-/config/train.yaml}"
ATTEMPTS="${DISCOVERY_ATTEMPTS:-30}"
PAUSE="${DISCOVERY_PAUSE_SEC:-2}"

log() { echo "[entrypoint rank=${RANK}] $*" >&2; }

find_master() {
  local sa=/var/run/secrets/kubernetes.io/serviceaccount
  local ns token selector
  ns="$(cat "${sa}/namespace")"
  token="$(cat "${sa}/token")"
  selector="batch.kubernetes.io/job-completion-index=0,job-name=${JOB_NAME}"
  curl -sS --fail --cacert "${sa}/ca.crt" -H "Authorization: Bearer ${token}" \
    "https://kubernetes.default.svc/api/v1/namespaces/${ns}/pods?labelSelector=${selector}" \
    | jq -r '[.items[] | select(.status.podIP != null)][0].status.podIP // empty'
}

if [[ "${RANK}" == "0" ]]; then
  : "${POD_IP:?rank 0 needs POD_IP (downward API status.podIP)}"
  export MASTER_ADDR="${POD_IP}"
else
  MASTER_ADDR=""
  for (( i = 1; i <= ATTEMPTS; i++ )); do
    MASTER_ADDR="$(find_master || true)"
    [[ -n "${MASTER_ADDR}" ]] && break
    log "rank-0 pod has no IP yet (${i}/${ATTEMPTS}); retrying in ${PAUSE}s"
    sleep "${PAUSE}"
  done
  if [[ -z "${MASTER_ADDR}" ]]; then
    log "giving up: could not discover the rank-0 pod"
    exit 1
  fi
  export MASTER_ADDR
fi

log "WORLD_SIZE=${WORLD_SIZE} LOCAL_RANK=${LOCAL_RANK} MASTER=${MASTER_ADDR}:${MASTER_PORT} config=${CONFIG}"
exec python -m llmtrain train --config "${CONFIG}" "$@"
