"""Operator surface: kind-gpu-sim.sh and the probe pod manifests.

kind / kubectl / docker do not exist in this image, so the script runs against PATH shims
that record every call; the assertions are the names, labels, taints, patches and manifests
the reference's users and CI depend on (Appendix A of SURVEY.md), plus the additions
(`create nvidia-real`, --gpus-per-worker=) and what they must NOT do (no fake capacity patch)."""
import os
import stat
import subprocess
import sys
import textwrap

import pytest
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = os.path.join(ROOT, "kind-gpu-sim.sh")

SHIM = r'''#!/bin/bash
# records "<tool> <args>" and fakes the few outputs the script parses
tool=$(basename "$0")
echo "$tool $*" >> "$SHIM_LOG"
case "$tool $*" in
  "kubectl get nodes -o jsonpath={range .items[*]}{.metadata.name}"*) printf 'c-control-plane\nc-worker\nc-worker2\n' ;;
  "kubectl get nodes -o jsonpath={range .items[*]}{.status.allocatable"*) printf '%s\n' $SHIM_ALLOCATABLE ;;
  "kubectl apply -f -") n=$(ls "$SHIM_DIR"/applied_*.yaml 2>/dev/null | wc -l); cat > "$SHIM_DIR/applied_$n.yaml" ;;
  "kind get nodes"*) printf 'c-control-plane\nc-worker\nc-worker2\n' ;;
  "kind get clusters") echo "${SHIM_CLUSTERS:-kind-gpu-sim}" ;;
  "docker inspect"*) echo false ;;
  "docker exec -i"*) cat > /dev/null ;;
  "nvidia-smi -L") for i in $(seq 0 $((${SHIM_HOST_GPUS:-8} - 1))); do echo "GPU $i: NVIDIA B200 (UUID: GPU-$i)"; done ;;
  "git clone"*) d="${@: -1}"; mkdir -p "$d/deployments/container"; echo "FROM redhat/ubi9-minimal" > "$d/deployments/container/Dockerfile"; echo "FROM alpine:3.21.3" > "$d/Dockerfile" ;;
esac
exit 0
'''


@pytest.fixture(params=[()], ids=["docker"])
def shims(tmp_path, request):
    bindir = tmp_path / "bin"
    bindir.mkdir()
    for tool in ("kind", "kubectl", "docker", "git", "nvidia-smi", "sleep", "pidof") + tuple(request.param):
        p = bindir / tool
        p.write_text(SHIM)
        p.chmod(p.stat().st_mode | stat.S_IEXEC)
    work = tmp_path / "work"
    work.mkdir()
    env = {"PATH": f"{bindir}:/usr/bin:/bin", "SHIM_LOG": str(tmp_path / "calls.log"), "SHIM_DIR": str(tmp_path),
           "HOME": str(tmp_path), "UID": "0"}

    def run(*args, **extra):
        (tmp_path / "calls.log").write_text("")
        for f in tmp_path.glob("applied_*.yaml"):
            f.unlink()
        r = subprocess.run(["/bin/bash", SCRIPT, *args], cwd=work, env={**env, **extra}, capture_output=True, text=True)
        calls = (tmp_path / "calls.log").read_text().splitlines()
        applied = [list(yaml.safe_load_all(f.read_text())) for f in sorted(tmp_path.glob("applied_*.yaml"))]
        return r, calls, [d for docs in applied for d in docs], work
    return run


@pytest.mark.parametrize("shims", [("podman", "systemctl")], indirect=True, ids=["podman"])
def test_podman_runtime_loads_images_as_archives(shims):
    # reference behaviour with podman: no registry push; tag to localhost/, save, `kind load image-archive`
    r, calls, applied, _ = shims("create", "nvidia")
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Using Podman as container runtime" in r.stdout
    assert any(c.startswith("podman build -t localhost:5000/nvidia-device-plugin:dev") for c in calls)
    assert "podman tag localhost:5000/nvidia-device-plugin:dev localhost/nvidia-device-plugin:dev" in calls
    assert "podman save localhost/nvidia-device-plugin:dev -o /tmp/image.tar" in calls
    assert "kind load image-archive /tmp/image.tar --name kind-gpu-sim" in calls
    assert not any(c.startswith(("podman push", "docker ")) for c in calls)
    ds = next(d for d in applied if d["kind"] == "DaemonSet")
    assert ds["spec"]["template"]["spec"]["containers"][0]["image"] == "localhost/nvidia-device-plugin:dev"
    r, calls, _, _ = shims("load", "--image-name=localhost/gpu-probe:dev")
    assert "podman save localhost/gpu-probe:dev -o /tmp/image.tar" in calls
    assert "kind load image-archive /tmp/image.tar --name kind-gpu-sim" in calls


def test_script_parses():
    assert subprocess.run(["/bin/bash", "-n", SCRIPT]).returncode == 0


def test_usage_and_unknown_type(shims):
    r, _, _, _ = shims()
    assert r.returncode == 1 and "Usage:" in r.stdout and "create [rocm|nvidia" in r.stdout and "|delete|load}" in r.stdout
    r, calls, _, _ = shims("create", "tpu")
    assert r.returncode == 1 and "Unknown GPU type: tpu" in r.stdout
    assert not any(c.startswith("kind create") for c in calls)


def test_create_nvidia_fake_keeps_the_reference_surface(shims):
    r, calls, applied, work = shims("create", "nvidia")
    assert r.returncode == 0, r.stderr
    assert " Simulated GPU Kind cluster is ready for 'nvidia'!" in r.stdout
    assert "kind create cluster --name kind-gpu-sim --config kind-config.yaml" in calls
    cfg = yaml.safe_load((work / "kind-config.yaml").read_text())
    assert [n["role"] for n in cfg["nodes"]] == ["control-plane", "worker", "worker"]
    assert all("extraMounts" not in n for n in cfg["nodes"])
    assert 'registry.mirrors."localhost:5000"' in cfg["containerdConfigPatches"][0]
    for node in ("c-worker", "c-worker2"):
        assert f"kubectl label node {node} hardware-type=gpu --overwrite" in calls
        assert f"kubectl label node {node} node-role.kubernetes.io/worker= --overwrite" in calls
        assert f"kubectl taint node {node} gpu=true:NoSchedule --overwrite" in calls
        assert f"kubectl label node {node} nvidia.com/gpu.present=true --overwrite" in calls
        patch = [c for c in calls if c.startswith(f"kubectl patch node {node} ")]
        assert len(patch) == 1 and "/status/capacity/nvidia.com~1gpu" in patch[0] and '"value":"2"' in patch[0]
        assert "--subresource=status" in patch[0]
    assert not any("c-control-plane" in c and ("label" in c or "taint" in c) for c in calls)
    assert any(c.startswith("docker run -d --restart=always -p 5000:5000 --name kind-registry "
                            "public.ecr.aws/docker/library/registry:2") for c in calls)
    assert "git checkout v0.18.2" in calls
    assert any(c.startswith("docker build -t localhost:5000/nvidia-device-plugin:dev") for c in calls)
    assert "docker push localhost:5000/nvidia-device-plugin:dev" in calls
    cm = next(d for d in applied if d["kind"] == "ConfigMap")
    assert cm["metadata"] == {"name": "local-registry-hosting", "namespace": "kube-public"}
    ds = next(d for d in applied if d["kind"] == "DaemonSet")
    assert ds["metadata"] == {"name": "nvidia-device-plugin-daemonset", "namespace": "kube-system"}
    tpl = ds["spec"]["template"]
    assert tpl["metadata"]["labels"] == {"app": "nvidia-device-plugin"} == ds["spec"]["selector"]["matchLabels"]
    assert tpl["spec"]["nodeSelector"] == {"hardware-type": "gpu"}
    assert tpl["spec"]["tolerations"] == [{"key": "gpu", "operator": "Equal", "value": "true", "effect": "NoSchedule"}]
    ctr = tpl["spec"]["containers"][0]
    assert ctr["name"] == "nvidia-device-plugin-ctr" and ctr["image"] == "localhost:5000/nvidia-device-plugin:dev"
    assert ctr["env"] == [{"name": "FAIL_ON_INIT_ERROR", "value": "false"}]
    assert ctr["volumeMounts"][0]["mountPath"] == "/var/lib/kubelet/device-plugins"
    assert any("kubectl wait --for=condition=Ready -n kube-system pod -l app=nvidia-device-plugin --timeout=60s" == c
               for c in calls)


def test_create_defaults_to_rocm_and_flags_are_honoured(shims):
    r, calls, applied, _ = shims("create", "--registry-port=5001", "--cluster-name=mycluster")
    assert r.returncode == 0, r.stderr
    assert "ready for 'rocm'!" in r.stdout
    assert "kind create cluster --name mycluster --config kind-config.yaml" in calls
    assert any("/status/capacity/amd.com~1gpu" in c for c in calls)
    assert "kubectl label node c-worker rocm.amd.com/gpu.present=true --overwrite" in calls
    ds = next(d for d in applied if d["kind"] == "DaemonSet")
    assert ds["metadata"]["name"] == "amdgpu-device-plugin-daemonset"
    ctr = ds["spec"]["template"]["spec"]["containers"][0]
    assert ctr["name"] == "amdgpu-dp-ds" and ctr["image"] == "localhost:5001/amdgpu-dp:dev"
    assert ctr["imagePullPolicy"] == "IfNotPresent"
    assert any(c.startswith("docker run -d --restart=always -p 5001:5000") for c in calls)


def test_create_nvidia_real_passes_gpus_through_and_never_fakes_capacity(shims):
    r, calls, applied, work = shims("create", "nvidia-real", SHIM_ALLOCATABLE="8 0")
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ready for 'nvidia-real'!" in r.stdout and "advertises 8 real nvidia.com/gpu" in r.stdout
    assert not any(c.startswith("kubectl patch node") for c in calls), "real mode must not patch fake capacity"
    cfg = yaml.safe_load((work / "kind-config.yaml").read_text())
    w1, w2 = [n for n in cfg["nodes"] if n["role"] == "worker"]
    assert [m["containerPath"] for m in w1["extraMounts"]] == [f"/var/run/nvidia-container-devices/{i}" for i in range(8)]
    assert "extraMounts" not in w2
    assert "kubectl label node c-worker nvidia.com/gpu.present=true --overwrite" in calls
    assert "kubectl label node c-worker2 nvidia.com/gpu.present=true --overwrite" not in calls
    assert "kubectl taint node c-worker2 gpu=true:NoSchedule --overwrite" in calls
    assert any(c.startswith("docker exec c-worker nvidia-ctk runtime configure --runtime=containerd") for c in calls)
    assert not any(c.startswith("docker exec c-worker2 nvidia-ctk") for c in calls)
    ds = next(d for d in applied if d["kind"] == "DaemonSet")
    assert ds["metadata"]["name"] == "nvidia-device-plugin-daemonset"


def test_nvidia_real_gpu_layouts(shims):
    r, _, _, work = shims("create", "nvidia-real", "--gpus-per-worker=4", SHIM_ALLOCATABLE="4 4")
    assert r.returncode == 0, r.stdout + r.stderr
    cfg = yaml.safe_load((work / "kind-config.yaml").read_text())
    w1, w2 = [n for n in cfg["nodes"] if n["role"] == "worker"]
    assert [m["containerPath"][-1] for m in w1["extraMounts"]] == list("0123")
    assert [m["containerPath"][-1] for m in w2["extraMounts"]] == list("4567")
    r, _, _, _ = shims("create", "nvidia-real", "--gpus-per-worker=8,1")
    assert r.returncode == 1 and "asks for 9 GPUs, the host has 8" in r.stderr
    r, _, _, _ = shims("create", "nvidia-real", "--gpus-per-worker=x")
    assert r.returncode == 1 and "not a number" in r.stderr
    r, _, _, _ = shims("create", "nvidia-real", SHIM_ALLOCATABLE="2 0")   # plugin reports too few
    assert r.returncode == 1 and "expected 8" in r.stderr


def test_plugin_image_override_skips_the_build(shims):
    r, calls, applied, _ = shims("create", "nvidia", "--plugin-image=nvcr.io/nvidia/k8s-device-plugin:v0.18.2")
    assert r.returncode == 0
    assert not any(c.startswith(("git clone", "docker build")) for c in calls)
    ds = next(d for d in applied if d["kind"] == "DaemonSet")
    assert ds["spec"]["template"]["spec"]["containers"][0]["image"] == "nvcr.io/nvidia/k8s-device-plugin:v0.18.2"


def test_delete_and_load(shims):
    r, calls, _, _ = shims("delete")
    assert r.returncode == 0 and "kind delete cluster --name kind-gpu-sim" in calls
    assert "docker stop kind-registry" in calls and "docker rm kind-registry" in calls
    r, calls, _, _ = shims("delete", "--cluster-name=other")
    assert "Kind cluster 'other' does not exist. Skipping delete." in r.stdout
    assert not any(c.startswith("kind delete") for c in calls)
    r, calls, _, _ = shims("load", "--image-name=localhost:5000/gpu-probe:dev")
    assert r.returncode == 0 and "kind load docker-image localhost:5000/gpu-probe:dev --name kind-gpu-sim" in calls


@pytest.mark.parametrize("name,gpus", [("probe-1gpu.yaml", 1), ("probe-8gpu.yaml", 8)])
def test_probe_pod_manifests_keep_the_scheduling_contract(name, gpus):
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import pod_runner
    pod = pod_runner.load_pod(os.path.join(ROOT, "pods", name))
    assert pod_runner.placement_contract(pod) == {"nodeSelector_ok": True, "toleration_ok": True}
    cmd, limit = pod_runner.container_of(pod)
    assert limit == gpus and cmd[0] == "/probe/bin/gpu-probe" and f"--expect-gpus={gpus}" in cmd
    assert "--bytes=4294967296" in cmd                      # BASELINE.json: 4 GiB arrays
    assert pod["spec"]["restartPolicy"] == "Never"
    assert pod["spec"]["tolerations"][0]["value"] == "true"  # a string, not the YAML boolean
    env = {e["name"]: e["value"] for e in pod["spec"]["containers"][0]["env"]}
    assert float(env["GPU_PROBE_MEASURED_HBM_GBS"]) > 0 and isinstance(env["GPU_PROBE_MEASURED_HBM_GBS"], str)


def test_pod_runner_maps_the_image_path_and_reports_exit_code():
    # without a GPU the container command must fail with the device-count exit code (3)
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("GPU present: covered by the gpu-marked pod test")
    except ImportError:
        pass
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "pod_runner.py"),
                        os.path.join(ROOT, "pods", "probe-1gpu.yaml")], capture_output=True, text=True)
    assert r.returncode == 3 and "PROBE FAIL rc=3" in r.stdout and "POD name=gpu-probe-1" in r.stdout
