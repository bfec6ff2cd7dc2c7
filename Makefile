# Developer entry points.  `make help` lists them.
PY ?= python
NPROC ?= 2
CONFIG ?= configs/presets/ddp_smoke.yaml
IMAGE ?= llmtrain-b200:latest
KIND_CLUSTER ?= llmtrain

.PHONY: help build test test-gpu sanitize conformance format lint typecheck bench bench-ref \
        train-ddp train-gpt-ddp train-b200 mlflow \
        k8s-cluster k8s-cluster-delete k8s-build k8s-train k8s-train-cpu k8s-logs k8s-clean k8s-full \
        k8s-e2e k8s-mlflow k8s-dashboard k8s-dashboard-token k8s-dashboard-proxy k8s-dashboard-delete

help:
	@grep -E '^[a-zA-Z0-9_-]+:.*?## ' $(MAKEFILE_LIST) | awk 'BEGIN{FS=":.*?## "}{printf "%-20s %s\n",$$1,$$2}'

build: ## compile csrc/**/*.cu for sm_100a into src/llmtrain/_native/
	$(PY) __graft_entry__.py

test: ## CPU test-suite
	$(PY) -m pytest tests -q -m "not gpu"

test-gpu: ## GPU test-suite (needs a B200)
	$(PY) -m pytest tests -q -m gpu

sanitize: ## compute-sanitizer memcheck + racecheck + synccheck over every native kernel (needs a B200)
	$(PY) bench/sanitize_probe.py
	scripts/sanitize.sh memcheck
	scripts/sanitize.sh racecheck
	scripts/sanitize.sh synccheck

conformance: ## run the reference's own tests against this package
	scripts/run_conformance.sh

format: ## rewrite sources in the house style
	ruff format src tests bench.py bench scripts

lint: ## style + static checks (what the pre-commit hooks run)
	ruff check src tests bench.py && ruff format --check src tests
	$(MAKE) typecheck

typecheck: ## mypy over the package (configuration in pyproject.toml; skipped when mypy is absent)
	@if $(PY) -c "import mypy" 2>/dev/null; then $(PY) -m mypy --config-file=pyproject.toml src/llmtrain; \
	 else echo "mypy is not installed: skipping"; fi

bench: ## headline benchmark on 1 GPU (torchrun for more: see bench.py)
	$(PY) bench.py --gpus 1 --steps 10 --warmup 3

bench-ref: ## same benchmark through the unmodified reference (baseline/_ref)
	$(PY) bench.py --impl reference --gpus 1 --steps 10 --warmup 3

train-ddp: ## CPU/gloo data-parallel smoke run
	$(PY) -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node $(NPROC) \
	  -m llmtrain train --config $(CONFIG)

train-gpt-ddp: ## CPU/gloo data-parallel GPT run (WikiText-2 preset; needs the dataset cache or network)
	$(PY) -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node $(NPROC) \
	  -m llmtrain train --config configs/presets/gpt_wikitext_ddp.yaml

train-b200: ## GPT-2 small on all local B200s over NVLink
	$(PY) -m torch.distributed.run --standalone --local-addr 127.0.0.1 --nproc-per-node 8 \
	  -m llmtrain train --config configs/presets/gpt2_small_b200_ddp.yaml

mlflow: ## MLflow UI over the local SQLite store
	mlflow ui --backend-store-uri sqlite:///./mlflow.db

k8s-cluster: ## local kind cluster (CPU plumbing path)
	mkdir -p runs mlflow-k8s && kind create cluster --name $(KIND_CLUSTER) --config k8s/kind-config.yaml

k8s-cluster-delete:
	kind delete cluster --name $(KIND_CLUSTER)

k8s-build: ## build the image and load it into kind
	docker build -t $(IMAGE) -f k8s/Dockerfile . && kind load docker-image $(IMAGE) --name $(KIND_CLUSTER)

k8s-train: ## 8 x B200 Indexed Job
	kubectl apply -f k8s/rbac.yaml -f k8s/storage.yaml -f k8s/configmap.yaml -f k8s/service.yaml -f k8s/job.yaml
	kubectl wait --for=condition=complete job/llmtrain --timeout=1800s

k8s-train-cpu: ## 2-pod CPU/gloo Indexed Job on kind
	kubectl apply -f k8s/rbac.yaml -f k8s/storage.yaml -f k8s/configmap-cpu.yaml -f k8s/service.yaml
	sed -e 's/completions: 8/completions: 2/' -e 's/parallelism: 8/parallelism: 2/' \
	    -e 's/value: "8"/value: "2"/' -e '/nvidia.com\/gpu/d' -e '/resources:/d' k8s/job.yaml | kubectl apply -f -
	kubectl wait --for=condition=complete job/llmtrain --timeout=300s

k8s-logs:
	kubectl logs -l app=llmtrain --prefix --tail=-1

k8s-clean:
	-kubectl delete job llmtrain llmtrain-node --ignore-not-found
	-kubectl delete -f k8s/service.yaml -f k8s/configmap.yaml -f k8s/storage.yaml -f k8s/rbac.yaml --ignore-not-found

k8s-e2e: ## full kind-based end-to-end check (CPU)
	k8s/test_e2e.sh

k8s-full: k8s-cluster k8s-build k8s-train-cpu k8s-logs ## cluster + image + 2-pod CPU job + logs in one go

k8s-mlflow: ## MLflow UI over the store the pods wrote through the hostPath mount
	mlflow ui --backend-store-uri sqlite:///mlflow-k8s/mlflow.db

# Kubernetes Dashboard (optional convenience for watching the Indexed Job; k8s/dashboard-admin.yaml
# grants the cluster-admin service account the token is minted for)
DASHBOARD_MANIFEST ?= https://raw.githubusercontent.com/kubernetes/dashboard/v2.7.0/aio/deploy/recommended.yaml
DASHBOARD_PATH := api/v1/namespaces/kubernetes-dashboard/services/https:kubernetes-dashboard:/proxy/

k8s-dashboard: ## deploy the dashboard, print a login token, start the proxy
	kubectl apply -f $(DASHBOARD_MANIFEST) -f k8s/dashboard-admin.yaml
	@$(MAKE) --no-print-directory k8s-dashboard-token
	@echo "open http://localhost:8001/$(DASHBOARD_PATH) and paste the token above"
	$(MAKE) --no-print-directory k8s-dashboard-proxy

k8s-dashboard-token: ## bearer token for the dashboard-admin service account
	@kubectl -n kubernetes-dashboard create token dashboard-admin

k8s-dashboard-proxy: ## kubectl proxy on :8001
	kubectl proxy

k8s-dashboard-delete:
	-kubectl delete -f k8s/dashboard-admin.yaml --ignore-not-found
	-kubectl delete -f $(DASHBOARD_MANIFEST) --ignore-not-found
