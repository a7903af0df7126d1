#!/usr/bin/env python
"""pod_runner.py -- run a probe pod manifest without a cluster (test and measurement aid).

kind, kubectl, docker and podman are absent from this image and from the GPU box, so the
last hop of the path -- kubelet starts the container command with the GPUs the device
plugin allocated -- is emulated here: read the manifest, take the container `command` and
the `nvidia.com/gpu` limit, expose exactly that many GPUs (CUDA_VISIBLE_DEVICES, which is
what the plugin's Allocate amounts to for the workload), map the image path /probe onto
the in-tree build, run it, and report launch -> first RESULT -> exit latency.

    python tools/pod_runner.py pods/probe-1gpu.yaml [--gpu-offset K] [--replicas R] [-- extra args]

--replicas R starts R copies at once on disjoint GPU sets (BASELINE.json configs[3]:
8 concurrent 1-GPU pods).  Exit code: the worst container exit code.

--beside NEIGHBOUR.yaml (BASELINE.json configs[4]: pods/vllm-cpu-pod.yaml co-scheduled with the
8-GPU probe): checks that the neighbour would land on the same tainted GPU workers (selector +
toleration, no GPU limit), then runs the probe pod twice -- alone, and beside a STAND-IN for the
neighbour's container: its image and model need a registry and the network, which this box does
not have, so the stand-in is one busy process per CPU the manifest binds
(VLLM_CPU_OMP_THREADS_BIND), pinned to that CPU, streaming through a few hundred MB of host
memory with no GPU visible.  Reports per-GPU triad GB/s with and without it.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROBE_ROOT = os.path.join(ROOT, "kind-gpu-sim_b200")


def measured_hbm_gbs():
    """The box's measured HBM copy bandwidth, when the driver left one (MEASURED_PEAKS.json)."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except Exception:
        return None


def load_pod(path: str) -> dict:
    with open(path) as f:
        pod = yaml.safe_load(f)
    if pod.get("kind") != "Pod":
        raise SystemExit(f"{path}: not a Pod manifest")
    return pod


def placement_contract(pod: dict) -> dict:
    """The scheduling fields every GPU pod of this repo must carry (reference pods/*.yaml)."""
    spec = pod["spec"]
    tol = spec.get("tolerations", [])
    return {
        "nodeSelector_ok": spec.get("nodeSelector", {}).get("hardware-type") == "gpu",
        "toleration_ok": any(t.get("key") == "gpu" and t.get("operator") == "Equal" and t.get("value") == "true"
                             and t.get("effect") == "NoSchedule" for t in tol),
    }


def neighbour_facts(pod: dict) -> dict:
    """What the co-scheduled CPU pod (reference pods/vllm-cpu-pod.yaml) asks of the cluster."""
    spec = pod["spec"]
    c = spec["containers"][0]
    env = {e["name"]: e.get("value") for e in c.get("env", [])}
    bind = str(env.get("VLLM_CPU_OMP_THREADS_BIND", ""))
    tol = [t for t in spec.get("tolerations", []) if t.get("key") == "gpu"]
    shm = [v for v in spec.get("volumes", []) if v.get("emptyDir", {}).get("medium") == "Memory"]
    return {
        "name": pod["metadata"]["name"], "image": c.get("image"),
        "gpu_limit": c.get("resources", {}).get("limits", {}).get("nvidia.com/gpu"),
        "nodeSelector_ok": spec.get("nodeSelector", {}).get("hardware-type") == "gpu",
        "toleration_key_effect_ok": any(t.get("operator") == "Equal" and t.get("effect") == "NoSchedule" for t in tol),
        # Toleration.value is a string in the Kubernetes API; an unquoted `true` is a YAML boolean
        "toleration_value": tol[0].get("value") if tol else None,
        "toleration_value_yaml_type": type(tol[0].get("value")).__name__ if tol else None,
        "cpus_bound": [int(x) for x in bind.split(",") if x.strip().isdigit()],
        "shm_size_limit": shm[0]["emptyDir"].get("sizeLimit") if shm else None,
        "privileged": bool(c.get("securityContext", {}).get("privileged")),
    }


STAND_IN = r"""
import os, sys
import numpy as np
cpu = int(sys.argv[1])
try:
    os.sched_setaffinity(0, {cpu})
except OSError:
    pass
n = 1 << 24                      # 3 x 128 MiB: misses every cache level, like a CPU-side KV cache walk
b, c = np.random.rand(n), np.random.rand(n)
a = np.empty_like(b)
print("ready", flush=True)
while True:
    np.multiply(c, 3.0, out=a)
    np.add(a, b, out=a)
"""


def start_stand_in(cpus):
    procs = []
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1", OPENBLAS_NUM_THREADS="1", MKL_NUM_THREADS="1")
    for cpu in cpus:
        p = subprocess.Popen([sys.executable, "-c", STAND_IN, str(cpu)], stdout=subprocess.PIPE, text=True, env=env)
        procs.append(p)
    for p in procs:
        p.stdout.readline()  # "ready": arrays touched, loop running
    return procs


def triad_gbs(lines):
    """per-GPU triad GB/s from the pod's RESULT lines (median launch, not best: the neighbour's effect is on the tail)"""
    out = {}
    for ln in lines:
        if ln.startswith("RESULT kind=triad ") and " gpu=all " not in ln:
            kv = dict(x.split("=", 1) for x in ln.split()[1:])
            out[kv["gpu"]] = round(float(kv["bytes"]) / (float(kv["ms_med"]) * 1e-3) / 1e9, 1)
    return out


def run_beside(args, pod, cmd, gpus):
    nb = neighbour_facts(load_pod(args.beside))
    report = {"probe_pod": pod["metadata"]["name"], "gpus_per_pod": gpus, "probe_placement": placement_contract(pod),
              "neighbour": nb,
              "neighbour_lands_on_gpu_workers": bool(nb["nodeSelector_ok"] and nb["toleration_key_effect_ok"]
                                                     and nb["gpu_limit"] is None),
              "stand_in": f"{len(nb['cpus_bound'])} busy processes pinned to CPUs {nb['cpus_bound']}, numpy triad over "
                          "3 x 128 MiB each, no GPU visible (the neighbour's image and model need the network)"}
    runs = {}
    for label in ("alone", "beside"):
        procs = start_stand_in(nb["cpus_bound"]) if label == "beside" else []
        try:
            out = {}
            run_one(cmd, gpus, args.gpu_offset, args.extra, out)
        finally:
            for p in procs:
                p.kill()
            for p in procs:
                p.wait()
        runs[label] = {"rc": out["rc"], "triad_GBps_per_gpu": triad_gbs(out["stdout"]),
                       "triad_pace_ns_per_gpu": {ln.split("gpu=")[1].split()[0]: int(ln.split("pace_ns=")[1].split()[0])
                                                 for ln in out["stdout"] if ln.startswith("PROBE INFO triad_calibrate")},
                       "launch_to_first_result_s": out["launch_to_first_result_s"],
                       "launch_to_exit_s": round(out["launch_to_exit_s"], 3),
                       "verdict": out["stdout"][-1] if out["stdout"] else "", "stderr_tail": out["stderr"][-400:]}
    report["runs"] = runs
    both = [g for g in runs["alone"]["triad_GBps_per_gpu"] if g in runs["beside"]["triad_GBps_per_gpu"]]
    report["triad_beside_over_alone"] = {g: round(runs["beside"]["triad_GBps_per_gpu"][g] / runs["alone"]["triad_GBps_per_gpu"][g], 4)
                                         for g in both}
    print(json.dumps(report, indent=1 if not args.json else None))
    sys.exit(max(runs["alone"]["rc"], runs["beside"]["rc"]))


POD_ENV: dict = {}


def container_of(pod: dict) -> tuple[list[str], int]:
    c = pod["spec"]["containers"][0]
    POD_ENV.clear()
    POD_ENV.update({e["name"]: str(e.get("value", "")) for e in c.get("env", []) if "name" in e})
    cmd = [str(x) for x in c.get("command", [])] + [str(x) for x in c.get("args", [])]
    gpus = int(c.get("resources", {}).get("limits", {}).get("nvidia.com/gpu", 0))
    return cmd, gpus


def run_one(cmd, gpus, first_gpu, extra, out):
    argv = [os.path.join(PROBE_ROOT, cmd[0][len("/probe/"):]) if cmd[0].startswith("/probe/") else cmd[0]] + cmd[1:] + extra
    env = dict(os.environ)
    env.update(POD_ENV)  # the container's own `env:` entries, as the kubelet would set them
    env["CUDA_VISIBLE_DEVICES"] = ",".join(str(first_gpu + i) for i in range(gpus))
    peak = measured_hbm_gbs()
    if peak:
        env["GPU_PROBE_MEASURED_HBM_GBS"] = str(peak)  # this box's own measurement beats the manifest's fleet figure
    t0 = time.time()
    p = subprocess.Popen(argv, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env)
    first_result, lines = None, []
    for line in p.stdout:
        if first_result is None and line.startswith("RESULT "):
            first_result = time.time() - t0
        lines.append(line.rstrip("\n"))
    err = p.stderr.read()
    rc = p.wait()
    out.update({"argv": argv, "visible": env["CUDA_VISIBLE_DEVICES"], "rc": rc, "stdout": lines, "stderr": err,
                "launch_to_first_result_s": first_result, "launch_to_exit_s": time.time() - t0})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("manifest")
    ap.add_argument("--gpu-offset", type=int, default=0)
    ap.add_argument("--replicas", type=int, default=1)
    ap.add_argument("--json", action="store_true", help="print one JSON summary instead of the pod logs")
    ap.add_argument("--beside", metavar="NEIGHBOUR.yaml", help="co-scheduling check: run the pod alone, then beside a stand-in for this CPU pod")
    argv = sys.argv[1:]
    extra = []
    if "--" in argv:  # everything after "--" is appended to the container command
        cut = argv.index("--")
        argv, extra = argv[:cut], argv[cut + 1:]
    args = ap.parse_args(argv)
    args.extra = extra
    pod = load_pod(args.manifest)
    cmd, gpus = container_of(pod)
    if gpus < 1:
        raise SystemExit("manifest requests no nvidia.com/gpu")
    if args.beside:
        run_beside(args, pod, cmd, gpus)
    outs = [dict() for _ in range(args.replicas)]
    thr = [threading.Thread(target=run_one, args=(cmd, gpus, args.gpu_offset + r * gpus, args.extra, outs[r]))
           for r in range(args.replicas)]
    for t in thr:
        t.start()
    for t in thr:
        t.join()
    worst = max(o["rc"] for o in outs)
    if args.json:
        print(json.dumps({"pod": pod["metadata"]["name"], "gpus_per_pod": gpus, "replicas": args.replicas,
                          "placement": placement_contract(pod), "runs": outs}))
    else:
        for r, o in enumerate(outs):
            for line in o["stdout"]:
                print(line if args.replicas == 1 else f"[pod {r} gpus={o['visible']}] {line}")
            sys.stderr.write(o["stderr"])
            print(f"POD name={pod['metadata']['name']} replica={r} gpus={o['visible']} rc={o['rc']} "
                  f"launch_to_first_result_s={o['launch_to_first_result_s']} launch_to_exit_s={o['launch_to_exit_s']:.3f}")
    sys.exit(worst)


if __name__ == "__main__":
    main()
