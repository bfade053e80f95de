"""Time the JSON -> columnar loader on an API-server-shaped PodList.

The synthetic documents carry the bulk a real `kubectl get pods -o json` has
(managedFields, volumes, env, probes, containerStatuses: about 3 KB per pod), of
which the accounting path reads a few dozen bytes.

    python tools/loader_timing.py [--pods 10000] [--nodes 100] [--reps 5] [--device]

Prints MB/s and pods/s of arena_load_json (no GPU needed); with --device also of
arena_load_json_device (PodList parsed on the GPU) followed by one `top node`.
"""
import argparse
import json
import random
import time


def make_pod(i, rng, n_nodes):
    name = f"job-{i // 4:05d}-worker-{i % 4}"
    node = f"node-{rng.randrange(n_nodes):04d}" if rng.random() < 0.95 else ""
    gpus = rng.choice([0, 0, 1, 1, 2, 4, 8])
    limits = {"cpu": "4", "memory": "16Gi"}
    if gpus:
        limits["nvidia.com/gpu"] = str(gpus)
    ctr = {
        "name": "main",
        "image": "registry.example.com/team/training-image:v1.2.3-cuda12",
        "command": ["python", "-m", "trainer.main", "--config", "/etc/config/train.yaml"],
        "env": [{"name": f"ENV_VAR_{k}", "value": f"value-{k}-{i}"} for k in range(8)],
        "resources": {"limits": limits, "requests": dict(limits)},
        "volumeMounts": [{"name": f"vol-{k}", "mountPath": f"/mnt/data{k}", "readOnly": k % 2 == 0} for k in range(3)],
        "livenessProbe": {"httpGet": {"path": "/healthz", "port": 8080, "scheme": "HTTP"},
                          "initialDelaySeconds": 30, "timeoutSeconds": 5, "periodSeconds": 10,
                          "successThreshold": 1, "failureThreshold": 3},
        "terminationMessagePath": "/dev/termination-log",
        "terminationMessagePolicy": "File",
        "imagePullPolicy": "IfNotPresent",
        "securityContext": {"capabilities": {"drop": ["ALL"]}, "runAsNonRoot": True},
    }
    side = {
        "name": "sidecar", "image": "registry.example.com/infra/log-shipper:2.0",
        "resources": {"limits": {"cpu": "100m", "memory": "128Mi"}, "requests": {"cpu": "50m", "memory": "64Mi"}},
        "terminationMessagePath": "/dev/termination-log", "terminationMessagePolicy": "File",
        "imagePullPolicy": "IfNotPresent",
    }
    phase = rng.choice(["Running"] * 6 + ["Pending", "Succeeded", "Failed"])
    ts = f"2024-03-{1 + i % 28:02d}T{i % 24:02d}:{i % 60:02d}:00Z"
    return {
        "apiVersion": "v1", "kind": "Pod",
        "metadata": {
            "name": name, "generateName": name[:-1], "namespace": "default",
            "selfLink": f"/api/v1/namespaces/default/pods/{name}",
            "uid": f"{i:08x}-1111-2222-3333-444455556666", "resourceVersion": str(1000000 + i),
            "creationTimestamp": ts,
            "labels": {"app": "tf-horovod", "release": f"job-{i // 4:05d}", "role": "worker",
                       "controller-uid": f"{i:08x}-aaaa-bbbb-cccc-ddddeeeeffff"},
            "annotations": {"kubernetes.io/psp": "restricted", "scheduler.alpha.kubernetes.io/critical-pod": "",
                            "prometheus.io/scrape": "true", "prometheus.io/port": "8080"},
            "ownerReferences": [{"apiVersion": "batch/v1", "kind": "Job", "name": name[:-9],
                                 "uid": f"{i:08x}-9999-8888-7777-666655554444", "controller": True,
                                 "blockOwnerDeletion": True}],
            "managedFields": [{"manager": "kube-controller-manager", "operation": "Update", "apiVersion": "v1",
                               "time": ts, "fieldsType": "FieldsV1",
                               "fieldsV1": {"f:metadata": {"f:labels": {".": {}, "f:app": {}, "f:release": {}},
                                                           "f:ownerReferences": {".": {}}},
                                            "f:spec": {"f:containers": {"k:{\"name\":\"main\"}": {
                                                ".": {}, "f:image": {}, "f:resources": {".": {}, "f:limits": {}}}}}}}],
        },
        "spec": {
            "volumes": [{"name": f"vol-{k}", "persistentVolumeClaim": {"claimName": f"claim-{k}-{i}"}} for k in range(3)],
            "containers": [ctr, side],
            "restartPolicy": "Never", "terminationGracePeriodSeconds": 30, "dnsPolicy": "ClusterFirst",
            "serviceAccountName": "default", "serviceAccount": "default", "nodeName": node,
            "securityContext": {}, "schedulerName": "default-scheduler",
            "tolerations": [{"key": "node.kubernetes.io/not-ready", "operator": "Exists", "effect": "NoExecute",
                             "tolerationSeconds": 300},
                            {"key": "node.kubernetes.io/unreachable", "operator": "Exists", "effect": "NoExecute",
                             "tolerationSeconds": 300}],
            "priority": 0, "enableServiceLinks": True,
        },
        "status": {
            "phase": phase,
            "conditions": [{"type": t, "status": "True", "lastProbeTime": None, "lastTransitionTime": ts}
                           for t in ("Initialized", "Ready", "ContainersReady", "PodScheduled")],
            "hostIP": f"10.0.{i % 250}.{1 + i % 200}", "podIP": f"172.16.{i % 250}.{1 + i % 200}",
            "startTime": ts,
            "containerStatuses": [{"name": c, "state": {"running": {"startedAt": ts}}, "lastState": {}, "ready": True,
                                   "restartCount": 0, "image": "registry.example.com/team/training-image:v1.2.3-cuda12",
                                   "imageID": "docker-pullable://registry.example.com/team/training-image@sha256:" + "ab" * 32,
                                   "containerID": "docker://" + "cd" * 32} for c in ("main", "sidecar")],
            "qosClass": "Burstable",
        },
    }


def make_node(n):
    return {"metadata": {"name": f"node-{n:04d}", "labels": {"kubernetes.io/hostname": f"node-{n:04d}"}},
            "status": {"capacity": {"cpu": "96", "memory": "1Ti", "nvidia.com/gpu": "8"},
                       "allocatable": {"cpu": "95", "memory": "1000Gi", "nvidia.com/gpu": "8"},
                       "addresses": [{"type": "InternalIP", "address": f"10.0.{n // 250}.{n % 250}"},
                                     {"type": "Hostname", "address": f"node-{n:04d}"}]}}


def make_docs(n_pods, n_nodes, seed=1):
    rng = random.Random(seed)
    pods = json.dumps({"kind": "PodList", "apiVersion": "v1", "metadata": {"resourceVersion": "1"},
                       "items": [make_pod(i, rng, n_nodes) for i in range(n_pods)]})
    nodes = json.dumps({"kind": "NodeList", "apiVersion": "v1", "items": [make_node(n) for n in range(n_nodes)]})
    return pods.encode(), nodes.encode()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pods", type=int, default=10000)
    ap.add_argument("--nodes", type=int, default=100)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--device", action="store_true")
    args = ap.parse_args()
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from arena_b200.snapshot import HostSnapshot

    pods, nodes = make_docs(args.pods, args.nodes)
    best = float("inf")
    for _ in range(args.reps):
        t0 = time.perf_counter()
        snap = HostSnapshot.from_json(pods, nodes)
        best = min(best, time.perf_counter() - t0)
        del snap
    line = {"podlist_bytes": len(pods), "bytes_per_pod": len(pods) / args.pods, "pods": args.pods,
            "host_threads": os.cpu_count(), "load_ms": best * 1e3, "MB_per_s": len(pods) / best / 1e6,
            "pods_per_s": args.pods / best}
    if args.device:
        import arena_b200
        from arena_b200 import DeviceSnapshot
        arena_b200.init(0)
        pinned = arena_b200.PinnedBuffer(len(pods)).fill(pods)
        for label, src in (("device", pods), ("device_pinned", pinned)):
            dbest, stats = float("inf"), None
            for _ in range(args.reps + 2):
                t0 = time.perf_counter()
                dev, st = DeviceSnapshot.from_json_device(src, nodes)
                dt = time.perf_counter() - t0
                if dt < dbest:
                    dbest, stats = dt, st
                t0 = time.perf_counter()
                arena_b200.top_node(dev)
                top_ms = (time.perf_counter() - t0) * 1e3
                dev.close()
            line.update({label + "_load_ms": dbest * 1e3, label + "_MB_per_s": len(pods) / dbest / 1e6,
                         label + "_pods_per_s": args.pods / dbest, label + "_stats": stats, "top_node_ms": top_ms})
    print(json.dumps(line))


if __name__ == "__main__":
    main()
