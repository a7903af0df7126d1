#!/bin/bash
# kind-gpu-sim.sh -- kind cluster whose workers advertise GPUs: simulated (rocm | nvidia) or,
# new here, the machine's real NVIDIA GPUs (nvidia-real) so that probe pods run real work.
#
# Operator surface kept from the reference (maryamtahhan/kind-gpu-sim, kind-gpu-sim.sh):
#   verbs `create [rocm|nvidia]`, `delete`, `load`            (reference :380-399, usage :364-367)
#   flags --registry-port= --cluster-name= --image-name=      (reference :31-43)
#   labels hardware-type=gpu, node-role.kubernetes.io/worker="", <vendor>/gpu.present=true,
#   taint gpu=true:NoSchedule, resources nvidia.com/gpu / amd.com/gpu, fake count "2"
#                                                              (reference :108-116)
#   DaemonSet / container names in kube-system                 (reference :252-271, :295-314)
#   ./kind-config.yaml, registry container kind-registry       (reference :68-69, :85-97)
# Additions: `create nvidia-real`, --gpus-per-worker=, --plugin-image=, --workers=.
# Written from scratch around one table of per-vendor facts; the reference is the spec for
# names and behaviour only.
set -e

REGISTRY_PORT=5000
REGISTRY_IMAGE=public.ecr.aws/docker/library/registry:2
REGISTRY_NAME=kind-registry
CLUSTER_NAME=kind-gpu-sim
LOAD_IMAGE_NAME=not-set
CONFIG_FILE=kind-config.yaml
GPUS_PER_WORKER=""        # nvidia-real: "N" (each worker) or "a,b,..." (per worker); default: all on worker 1
PLUGIN_IMAGE=""           # use this device-plugin image instead of building one
WORKERS=2
FAKE_GPU_COUNT=2
NVIDIA_PLUGIN_REPO=https://github.com/NVIDIA/k8s-device-plugin.git
NVIDIA_PLUGIN_TAG=v0.18.2
ROCM_PLUGIN_REPO=https://github.com/RadeonOpenCompute/k8s-device-plugin.git

for arg in "$@"; do
  case "$arg" in
    --registry-port=*)   REGISTRY_PORT="${arg#*=}" ;;
    --cluster-name=*)    CLUSTER_NAME="${arg#*=}" ;;
    --image-name=*)      LOAD_IMAGE_NAME="${arg#*=}" ;;
    --gpus-per-worker=*) GPUS_PER_WORKER="${arg#*=}" ;;
    --plugin-image=*)    PLUGIN_IMAGE="${arg#*=}" ;;
    --workers=*)         WORKERS="${arg#*=}" ;;
  esac
done

die() { echo "ERROR: $*" >&2; exit 1; }

usage() {
  echo "Usage: $0 {create [rocm|nvidia|nvidia-real]|delete|load}"
  exit 1
}

# ---------------------------------------------------------------- platform and runtime
case "$(uname -s)" in Darwin) ON_MACOS=true ;; *) ON_MACOS=false ;; esac
if command -v gsed >/dev/null 2>&1; then SED=gsed; else SED=sed; fi
sed_in_place() {  # BSD sed wants an explicit (empty) backup suffix
  if [ "$ON_MACOS" = true ] && [ "$SED" = sed ]; then sed -i '' "$@"; else "$SED" -i "$@"; fi
}

pick_runtime() {
  if command -v podman >/dev/null 2>&1; then
    echo "Using Podman as container runtime"
    CONTAINER_RUNTIME=podman
    export KIND_EXPERIMENTAL_PROVIDER=podman
    export DOCKER_HOST="unix:///run/user/$UID/podman/podman.sock"
    [ "$ON_MACOS" = true ] || systemctl --user enable --now podman.socket || true
  elif command -v docker >/dev/null 2>&1; then
    echo "Using Docker as container runtime"
    CONTAINER_RUNTIME=docker
  else
    die "Neither Docker nor Podman is installed."
  fi
}
cr() { "$CONTAINER_RUNTIME" "$@"; }

# ---------------------------------------------------------------- per-vendor facts
# vendor_facts <type> sets: RESOURCE LABEL PLUGIN_NAME PLUGIN_APP PLUGIN_CTR PLUGIN_TAGGED REAL
vendor_facts() {
  REAL=false
  case "$1" in
    rocm)
      RESOURCE=amd.com/gpu LABEL=rocm.amd.com/gpu.present PLUGIN_TAGGED=amdgpu-dp:dev
      PLUGIN_NAME=amdgpu-device-plugin-daemonset PLUGIN_APP=amdgpu-device-plugin PLUGIN_CTR=amdgpu-dp-ds ;;
    nvidia|nvidia-real)
      RESOURCE=nvidia.com/gpu LABEL=nvidia.com/gpu.present PLUGIN_TAGGED=nvidia-device-plugin:dev
      PLUGIN_NAME=nvidia-device-plugin-daemonset PLUGIN_APP=nvidia-device-plugin PLUGIN_CTR=nvidia-device-plugin-ctr
      [ "$1" = nvidia-real ] && REAL=true ;;
    *) echo " Unknown GPU type: $1"; exit 1 ;;
  esac
  return 0
}

# ---------------------------------------------------------------- registry
start_local_registry() {
  echo "Starting local registry on port ${REGISTRY_PORT}..."
  if [ "$(cr inspect -f '{{.State.Running}}' "$REGISTRY_NAME" 2>/dev/null || echo false)" = true ]; then
    echo "Registry '${REGISTRY_NAME}' already running."
  else
    cr run -d --restart=always -p "${REGISTRY_PORT}:5000" --name "$REGISTRY_NAME" "$REGISTRY_IMAGE"
  fi
  echo "Ensuring the registry is connected to the Kind network..."
  cr network connect kind "$REGISTRY_NAME" 2>/dev/null || true
}

apply_local_registry_configmap() {
  kubectl apply -f - <<EOF
apiVersion: v1
kind: ConfigMap
metadata:
  name: local-registry-hosting
  namespace: kube-public
data:
  localRegistryHosting.v1: |
    host: "localhost:${REGISTRY_PORT}"
    help: "https://kind.sigs.k8s.io/docs/user/local-registry/"
EOF
}

# ---------------------------------------------------------------- real GPUs: who gets which
# Fills GPU_SHARE[w] = number of GPUs of worker w (0-based) and GPU_FIRST[w] = its first index.
plan_real_gpus() {
  command -v nvidia-smi >/dev/null 2>&1 || die "nvidia-real needs nvidia-smi on the host"
  HOST_GPUS=$(nvidia-smi -L | grep -c '^GPU ' || true)
  [ "$HOST_GPUS" -ge 1 ] || die "nvidia-real: nvidia-smi lists no GPU"
  GPU_SHARE=() GPU_FIRST=()
  local spec="$GPUS_PER_WORKER" w next=0 share
  [ -n "$spec" ] || spec="$HOST_GPUS"   # default: every GPU on the first worker (room for the 8-GPU pod)
  for ((w = 0; w < WORKERS; w++)); do
    case "$spec" in
      *,*) share=$(echo "$spec" | cut -d, -f$((w + 1))); share=${share:-0} ;;
      *)   if [ -n "$GPUS_PER_WORKER" ]; then share=$spec; else share=$([ "$w" = 0 ] && echo "$spec" || echo 0); fi ;;
    esac
    case "$share" in ''|*[!0-9]*) die "--gpus-per-worker: '$share' is not a number" ;; esac
    GPU_SHARE[w]=$share GPU_FIRST[w]=$next
    next=$((next + share))
  done
  [ "$next" -le "$HOST_GPUS" ] || die "--gpus-per-worker asks for $next GPUs, the host has $HOST_GPUS"
  REAL_TOTAL=$next
}

# ---------------------------------------------------------------- kind config
generate_kind_config() {
  rm -f "$CONFIG_FILE"
  {
    cat <<EOF
kind: Cluster
apiVersion: kind.x-k8s.io/v1alpha4
containerdConfigPatches:
  - |
    [plugins."io.containerd.grpc.v1.cri".registry.mirrors."localhost:${REGISTRY_PORT}"]
      endpoint = ["http://${REGISTRY_NAME}:5000"]
nodes:
  - role: control-plane
EOF
    local w g
    for ((w = 0; w < WORKERS; w++)); do
      echo "  - role: worker"
      if [ "$REAL" = true ] && [ "${GPU_SHARE[w]}" -gt 0 ]; then
        # GPU passthrough: with `accept-nvidia-visible-devices-as-volume-mounts = true` in the host's
        # nvidia-container-runtime config, a mount under /var/run/nvidia-container-devices/<index>
        # makes the runtime inject that GPU into the node container.
        echo "    extraMounts:"
        for ((g = GPU_FIRST[w]; g < GPU_FIRST[w] + GPU_SHARE[w]; g++)); do
          echo "      - hostPath: /dev/null"
          echo "        containerPath: /var/run/nvidia-container-devices/${g}"
        done
      fi
    done
  } > "$CONFIG_FILE"
}

worker_nodes() {
  kubectl get nodes -o jsonpath='{range .items[*]}{.metadata.name}{"\n"}{end}' | grep -v control-plane
}

# Inside a worker that holds real GPUs: container toolkit + containerd's nvidia runtime as default.
prepare_real_worker() {
  local node="$1"
  cr exec "$node" bash -c 'command -v nvidia-ctk >/dev/null 2>&1 || {
      apt-get update -qq && apt-get install -y -qq gpg curl >/dev/null &&
      curl -fsSL https://nvidia.github.io/libnvidia-container/gpgkey | gpg --dearmor -o /usr/share/keyrings/nvidia-container-toolkit-keyring.gpg &&
      curl -fsSL https://nvidia.github.io/libnvidia-container/stable/deb/nvidia-container-toolkit.list |
        sed "s#deb https://#deb [signed-by=/usr/share/keyrings/nvidia-container-toolkit-keyring.gpg] https://#g" > /etc/apt/sources.list.d/nvidia-container-toolkit.list &&
      apt-get update -qq && apt-get install -y -qq nvidia-container-toolkit >/dev/null; }'
  # the node container inherits a masked /proc/driver/nvidia from the host runtime; the toolkit
  # inside the node needs the real one to enumerate the GPUs it was given
  cr exec "$node" umount -R /proc/driver/nvidia 2>/dev/null || true
  cr exec "$node" nvidia-ctk runtime configure --runtime=containerd --set-as-default --cdi.enabled
  cr exec "$node" systemctl restart containerd
  cr exec "$node" nvidia-smi -L
}

create_kind_cluster() {
  local gpu_type="$1" node w=0
  vendor_facts "$gpu_type"
  [ "$REAL" = true ] && plan_real_gpus
  generate_kind_config
  kind create cluster --name "$CLUSTER_NAME" --config "$CONFIG_FILE"
  cr network connect kind "$REGISTRY_NAME" || true

  for node in $(worker_nodes); do
    kubectl label node "$node" hardware-type=gpu --overwrite
    kubectl label node "$node" node-role.kubernetes.io/worker="" --overwrite
    kubectl taint node "$node" gpu=true:NoSchedule --overwrite
    if [ "$REAL" = true ]; then
      # no capacity patch: the device plugin reports what the node really holds
      if [ "${GPU_SHARE[w]:-0}" -gt 0 ]; then
        kubectl label node "$node" "${LABEL}=true" --overwrite
        prepare_real_worker "$node"
      fi
    else
      kubectl label node "$node" "${LABEL}=true" --overwrite
      kubectl patch node "$node" --type=json --subresource=status \
        -p="[{\"op\": \"add\", \"path\": \"/status/capacity/${RESOURCE/\//~1}\", \"value\":\"${FAKE_GPU_COUNT}\"}]"
    fi
    w=$((w + 1))
  done

  local pid_cmd=pidof
  [ "$ON_MACOS" = true ] && pid_cmd=pgrep
  for node in $(kind get nodes --name "$CLUSTER_NAME"); do
    cr exec "$node" mkdir -p "/etc/containerd/certs.d/localhost:${REGISTRY_PORT}"
    printf '[host."http://%s:5000"]\n  capabilities = ["pull", "resolve"]\n' "$REGISTRY_NAME" |
      cr exec -i "$node" tee "/etc/containerd/certs.d/localhost:${REGISTRY_PORT}/hosts.toml" >/dev/null
    cr exec "$node" kill -SIGHUP $($pid_cmd containerd) 2>/dev/null || echo "Warning: could not reload containerd on $node"
  done
}

# ---------------------------------------------------------------- device plugin image
# Registries the plugin Dockerfiles pull from that need no login.
retarget_base_images() {
  local file="$1"; shift
  local pair
  for pair in "$@"; do sed_in_place "s|^FROM ${pair%%=>*}|FROM ${pair##*=>}|" "$file"; done
}

ship_image() {  # registry push with docker; archive load with podman (no insecure-registry push)
  local tagged="$1"
  if [ "$CONTAINER_RUNTIME" = docker ]; then
    cr push "localhost:${REGISTRY_PORT}/${tagged}"
  else
    cr tag "localhost:${REGISTRY_PORT}/${tagged}" "localhost/${tagged}"
    cr save "localhost/${tagged}" -o /tmp/image.tar
    kind load image-archive /tmp/image.tar --name "$CLUSTER_NAME"
    rm -f /tmp/image.tar
  fi
}

build_and_push_images() {
  vendor_facts "$1"
  if [ -n "$PLUGIN_IMAGE" ]; then
    echo " Using device plugin image ${PLUGIN_IMAGE} (no local build)"
    return
  fi
  local ubi=registry.access.redhat.com/ubi9/ubi-minimal ecr=public.ecr.aws/docker/library
  if [ "$RESOURCE" = nvidia.com/gpu ]; then
    echo " Building NVIDIA device plugin locally..."
    [ -d k8s-device-plugin-nvidia ] || git clone "$NVIDIA_PLUGIN_REPO" k8s-device-plugin-nvidia
    ( cd k8s-device-plugin-nvidia
      git checkout "$NVIDIA_PLUGIN_TAG"
      if [ "$CONTAINER_RUNTIME" = podman ]; then
        retarget_base_images deployments/container/Dockerfile "redhat/ubi9-minimal=>$ubi" \
          "public.ecr.aws/ubi9/ubi-minimal=>$ubi" "registry.access.redhat.com/ubi9/ubi9-minimal=>$ubi"
        grep FROM deployments/container/Dockerfile
        export BUILDAH_FORMAT=docker
      fi
      cr build -t "localhost:${REGISTRY_PORT}/${PLUGIN_TAGGED}" -f deployments/container/Dockerfile . )
  else
    echo " Building ROCm plugin images locally..."
    [ -d k8s-device-plugin-rocm ] || git clone "$ROCM_PLUGIN_REPO" k8s-device-plugin-rocm
    ( cd k8s-device-plugin-rocm
      retarget_base_images Dockerfile "alpine:3.21.3=>$ecr/alpine:3.21.3" \
        "docker.io/golang:1.23.6-alpine3.21=>$ecr/golang:1.23.6-alpine3.21" \
        "golang:1.23.6-alpine3.21=>$ecr/golang:1.23.6-alpine3.21"
      cr build -t "localhost:${REGISTRY_PORT}/${PLUGIN_TAGGED}" -f Dockerfile . )
  fi
  ship_image "$PLUGIN_TAGGED"
}

# ---------------------------------------------------------------- device plugin DaemonSet
deploy_device_plugin() {
  vendor_facts "$1"
  local image="${PLUGIN_IMAGE:-localhost:${REGISTRY_PORT}/${PLUGIN_TAGGED}}"
  [ -z "$PLUGIN_IMAGE" ] && [ "$CONTAINER_RUNTIME" = podman ] && image="localhost/${PLUGIN_TAGGED}"
  {
    cat <<EOF
apiVersion: apps/v1
kind: DaemonSet
metadata:
  name: ${PLUGIN_NAME}
  namespace: kube-system
spec:
  selector:
    matchLabels:
      app: ${PLUGIN_APP}
  template:
    metadata:
      labels:
        app: ${PLUGIN_APP}
    spec:
      nodeSelector:
        hardware-type: gpu
      tolerations:
        - key: gpu
          operator: Equal
          value: "true"
          effect: NoSchedule
      containers:
        - name: ${PLUGIN_CTR}
          image: ${image}
          securityContext:
            privileged: true
EOF
    if [ "$RESOURCE" = amd.com/gpu ]; then
      echo "          imagePullPolicy: IfNotPresent"
    else
      # Fake mode has no NVML, so the plugin must idle instead of crashing.  Real mode keeps the same
      # setting because a worker may hold zero GPUs; readiness is judged by allocatable below.
      cat <<EOF
          env:
            - name: FAIL_ON_INIT_ERROR
              value: "false"
          volumeMounts:
            - name: device-plugin
              mountPath: /var/lib/kubelet/device-plugins
      volumes:
        - name: device-plugin
          hostPath:
            path: /var/lib/kubelet/device-plugins
            type: DirectoryOrCreate
EOF
    fi
  } | kubectl apply -f -

  sleep 5
  kubectl wait --for=condition=Ready -n kube-system pod -l "app=${PLUGIN_APP}" --timeout=60s || {
    case "$RESOURCE" in amd.com/gpu) echo >&2 "ERROR: ROCm plugin pods not ready in time" ;;
                        *) echo >&2 "ERROR: NVIDIA plugin pods not ready in time" ;; esac
    exit 1
  }
  [ "$REAL" = true ] && wait_for_real_capacity
  return 0
}

# Real mode only: the kubelet must end up advertising exactly the GPUs handed to the workers.
wait_for_real_capacity() {
  local tries total
  for ((tries = 0; tries < 30; tries++)); do
    total=$(kubectl get nodes -o jsonpath='{range .items[*]}{.status.allocatable.nvidia\.com/gpu}{"\n"}{end}' |
            awk '{s += $1} END {print s + 0}')
    if [ "$total" = "$REAL_TOTAL" ]; then
      echo " Device plugin advertises ${total} real nvidia.com/gpu"
      return 0
    fi
    sleep 2
  done
  die "device plugin advertises ${total} nvidia.com/gpu, expected ${REAL_TOTAL}"
}

# ---------------------------------------------------------------- delete / load
delete_cluster() {
  if kind get clusters | grep -qx "$CLUSTER_NAME"; then
    echo "Deleting kind cluster '${CLUSTER_NAME}'..."
    kind delete cluster --name "$CLUSTER_NAME"
  else
    echo "Kind cluster '${CLUSTER_NAME}' does not exist. Skipping delete."
  fi
}

delete_registry() {
  echo "Stopping ${REGISTRY_NAME} (if running)..."
  if cr ps -q -f "name=^/${REGISTRY_NAME}$" >/dev/null 2>&1; then
    cr stop "$REGISTRY_NAME" || echo "Warning: Failed to stop ${REGISTRY_NAME}"
  else
    echo "No running container named '${REGISTRY_NAME}' to stop."
  fi
  echo "Removing ${REGISTRY_NAME} (if exists)..."
  if cr ps -aq -f "name=^/${REGISTRY_NAME}$" >/dev/null 2>&1; then
    cr rm "$REGISTRY_NAME" || echo "Warning: Failed to remove ${REGISTRY_NAME}"
  else
    echo "No container named '${REGISTRY_NAME}' to remove."
  fi
}

load_image() {
  if [ "$CONTAINER_RUNTIME" = docker ]; then
    echo "Running: load docker-image ${LOAD_IMAGE_NAME} --name ${CLUSTER_NAME}"
    kind load docker-image "$LOAD_IMAGE_NAME" --name "$CLUSTER_NAME"
  else
    cr save "$LOAD_IMAGE_NAME" -o /tmp/image.tar
    kind load image-archive /tmp/image.tar --name "$CLUSTER_NAME"
    rm -f /tmp/image.tar
  fi
}

# ---------------------------------------------------------------- verbs
case "$1" in
  create)
    gpu_type=${2:-rocm}
    case "$gpu_type" in --*) gpu_type=rocm ;; esac
    vendor_facts "$gpu_type"
    pick_runtime
    start_local_registry
    create_kind_cluster "$gpu_type"
    apply_local_registry_configmap
    build_and_push_images "$gpu_type"
    deploy_device_plugin "$gpu_type"
    echo " Simulated GPU Kind cluster is ready for '${gpu_type}'!"
    ;;
  delete)
    pick_runtime
    delete_cluster
    delete_registry
    ;;
  load)
    pick_runtime
    load_image
    ;;
  *)
    usage
    ;;
esac
