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


def container_of(pod: dict) -> tuple[list[str], int]:
    c = pod["spec"]["containers"][0]
    cmd = [str(x) for x in c.get("command", [])] + [str(x) for x in c.get("args", [])]
    gpus = int(c.get("resources", {}).get("limits", {}).get("nvidia.com/gpu", 0))
    return cmd, gpus


def run_one(cmd, gpus, first_gpu, extra, out):
    argv = [os.path.join(PROBE_ROOT, cmd[0][len("/probe/"):]) if cmd[0].startswith("/probe/") else cmd[0]] + cmd[1:] + extra
    env = dict(os.environ)
    env["CUDA_VISIBLE_DEVICES"] = ",".join(str(first_gpu + i) for i in range(gpus))
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
