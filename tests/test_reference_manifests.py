"""The reference's own pod manifests ship unchanged ("existing pods/*.yaml
unchanged"; BASELINE.json configs[0] and configs[4] run from them).

They are configuration a user of the reference already applies with `kubectl create -f pods/...`
(/root/reference/.github/workflows/nvidia-ci.yaml:32-39, Readme.md:75), so they are kept byte for
byte: compared with /root/reference/pods when that tree is mounted, else with the digests below."""
import hashlib
import os
import subprocess
import sys

import pytest
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PODS = os.path.join(ROOT, "pods")
REF_PODS = "/root/reference/pods"
sys.path.insert(0, os.path.join(ROOT, "tools"))
import pod_runner  # noqa: E402

SHA256 = {
    "nvidia-gpu-test-pod.yaml": "8e0f220f8949280ba95e518a1b9b211e442a6c52d15b8e21c340367ee3687888",
    "rocm-gpu-test-pod.yaml": "1d626aab0b35eba3dff043c3d1aafe19b0cd0a015a8239721f27d49bbb3fc735",
    "triton-pod.yaml": "4b26cf2bf6f7c52e143fa0a3c78a4709b6845489322c0b631f816a0eb48f96df",
    "vllm-cpu-pod.yaml": "8153552c3b0473b878d5d738e806f4f333080cd12a9b6308f3152f35d4c22f23",
}


def test_pods_directory_holds_the_four_reference_manifests_and_the_two_probe_pods():
    assert sorted(os.listdir(PODS)) == sorted(list(SHA256) + ["probe-1gpu.yaml", "probe-8gpu.yaml"])


@pytest.mark.parametrize("name", sorted(SHA256))
def test_reference_manifest_is_byte_identical(name):
    data = open(os.path.join(PODS, name), "rb").read()
    assert hashlib.sha256(data).hexdigest() == SHA256[name]
    ref = os.path.join(REF_PODS, name)
    if os.path.exists(ref):  # in the build container; the GPU box has no /root/reference
        assert data == open(ref, "rb").read()


def test_config1_pod_is_the_echo_workload_on_one_fake_gpu():
    # /root/reference/pods/nvidia-gpu-test-pod.yaml:9-19 -- the slot the probe pods fill
    pod = pod_runner.load_pod(os.path.join(PODS, "nvidia-gpu-test-pod.yaml"))
    assert pod["metadata"]["name"] == "nvidia-gpu-test"
    cmd, gpus = pod_runner.container_of(pod)
    assert gpus == 1 and cmd == ["sh", "-c", "echo Hello from fake NVIDIA GPU node && sleep 3600"]
    assert pod_runner.placement_contract(pod) == {"nodeSelector_ok": True, "toleration_ok": True}
    # the echo is the whole workload: running it as the kubelet would prints the README's line (Readme.md:82)
    out = subprocess.run(["sh", "-c", cmd[2].replace("sleep 3600", "true")], capture_output=True, text=True)
    assert out.stdout.strip() == "Hello from fake NVIDIA GPU node"


def test_config5_neighbour_shares_the_probe_pods_placement_and_asks_for_no_gpu():
    # /root/reference/pods/vllm-cpu-pod.yaml:26-37
    nb = pod_runner.neighbour_facts(pod_runner.load_pod(os.path.join(PODS, "vllm-cpu-pod.yaml")))
    assert nb["name"] == "vllm-cpu-pod" and nb["gpu_limit"] is None
    assert nb["nodeSelector_ok"] and nb["toleration_key_effect_ok"]
    assert nb["cpus_bound"] == [0, 1, 2, 3]          # VLLM_CPU_OMP_THREADS_BIND, :14-15
    assert nb["shm_size_limit"] == "4Gi" and nb["privileged"]
    probe = pod_runner.load_pod(os.path.join(PODS, "probe-8gpu.yaml"))
    assert probe["spec"]["nodeSelector"] == yaml.safe_load(open(os.path.join(PODS, "vllm-cpu-pod.yaml")))["spec"]["nodeSelector"]


def test_config5_neighbour_toleration_value_is_a_yaml_boolean_and_that_is_documented():
    """vllm-cpu-pod.yaml:31 writes `value: true` unquoted.  YAML makes that a boolean; the
    Kubernetes Toleration.value field is a string, so a client that validates strictly refuses
    the manifest (the other three manifests quote it).  Kept as the reference has it -- the file
    is theirs -- and called out in INTEGRATION.md for anyone who applies it."""
    nb = pod_runner.neighbour_facts(pod_runner.load_pod(os.path.join(PODS, "vllm-cpu-pod.yaml")))
    assert nb["toleration_value"] is True and nb["toleration_value_yaml_type"] == "bool"
    for name in ("nvidia-gpu-test-pod.yaml", "rocm-gpu-test-pod.yaml", "triton-pod.yaml", "probe-1gpu.yaml", "probe-8gpu.yaml"):
        tol = pod_runner.load_pod(os.path.join(PODS, name))["spec"]["tolerations"][0]
        assert tol["value"] == "true" and isinstance(tol["value"], str), name
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    assert "vllm-cpu-pod.yaml" in text and "value: true" in text


def test_stand_in_neighbour_starts_pinned_and_stops():
    procs = pod_runner.start_stand_in([0])
    try:
        assert procs[0].poll() is None
        if hasattr(os, "sched_getaffinity"):
            assert os.sched_getaffinity(procs[0].pid) == {0}
    finally:
        for p in procs:
            p.kill()
            p.wait()


def test_pod_runner_beside_reports_both_runs_even_when_the_probe_cannot_start():
    """Config 5's runner on a machine without a GPU: the placement facts are still reported, both probe runs fail
    with the device-count exit code (no CPU path), and that code is the runner's own."""
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("GPU present: the measured form is profiles/config5_r2.json")
    except ImportError:
        pass
    import json
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "pod_runner.py"), os.path.join(PODS, "probe-8gpu.yaml"),
                        "--beside", os.path.join(PODS, "vllm-cpu-pod.yaml"), "--json", "--", "--no-sweep", "--no-p2p",
                        "--no-sharded"], capture_output=True, text=True, timeout=300)
    out = json.loads(r.stdout)
    assert r.returncode == 3 and out["runs"]["alone"]["rc"] == 3 and out["runs"]["beside"]["rc"] == 3
    assert out["neighbour_lands_on_gpu_workers"] is True and out["neighbour"]["cpus_bound"] == [0, 1, 2, 3]
    assert out["probe_placement"] == {"nodeSelector_ok": True, "toleration_ok": True}
    assert out["runs"]["alone"]["triad_GBps_per_gpu"] == {} and "PROBE FAIL rc=3" in out["runs"]["alone"]["verdict"]
