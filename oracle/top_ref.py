"""ORACLE (test infrastructure only; see oracle/quantity.py header).

Pure-Python restatement of the reference's `arena top node` / `arena top job`
from the JSON fixtures to the printed text, following the Go structure line by
line (nested loops and all -- small inputs only).  Paths relative to
/root/reference:

  gpu_in_container / _deprecated   cmd/arena/commands/gpu.go:61-79
  gpu_in_pod / gpu_in_active_pod   gpu.go:39-59
  gpu_in_node / _deprecated        gpu.go:18-37
  gpu_pods                         gpu.go:8-15
  acquire_all_active_pods          pod_helper.go:15-31 (+ fields/selector.go:118-120)
  get_all_node_infos               top_node.go:69-97   (nested-loop join)
  caculate_node_gpu                top_node.go:232-242
  display_top_node_summary         top_node.go:107-157
  display_top_node_details         top_node.go:159-229
  list_release_map                 util/helm/helm.go:164-195
  MPI / Standalone / TF trainers   trainer_mpi.go:91-119, 222-332;
                                   trainer_standalone.go:71-99, 182-273;
                                   trainer_tensorflow.go:69-86, 135-165, 215-245, 315-439
  JobInfo                          job_info.go:41-108
  ordering                         trainer.go:25-37, 68-75 (sort.Sort is unstable: ties
                                   are broken by name here)
  display_training_job_list        list.go:82-133
  short_human_duration             util/duration.go:10-27
  TabWriter                        Go text/tabwriter (standard library, NOT in the tree:
                                   "parity unpinned" except by the doc samples,
                                   docs/userguide/1-tfjob-standalone.md:7-17, 73-98)
"""

from __future__ import annotations

import calendar
import json
import re
from typing import Dict, List, Optional, Tuple

from . import quantity as Q

NVIDIA = "nvidia.com/gpu"                       # const.go:7
DEPRECATED = "alpha.kubernetes.io/nvidia-gpu"   # const.go:9
MASTER_LABEL = "node-role.kubernetes.io/master"  # const.go:11
ZERO_TIME = -62135596800


class RawNum(str):
    """A JSON number kept as its raw token text (Quantity.UnmarshalJSON sees raw bytes)."""


def loads(text) -> dict:
    if isinstance(text, bytes):
        text = text.decode("utf-8")
    return json.loads(text, parse_int=RawNum, parse_float=RawNum)


def wrap(x: int) -> int:
    return Q.wrap_i64(x)


class DecodeError(ValueError):
    pass


def _raw_token(v) -> str:
    if v is None:
        return "null"
    if isinstance(v, RawNum):
        return str(v)
    if isinstance(v, bool):
        return "true" if v else "false"
    if isinstance(v, str):
        return '"' + v + '"'
    return json.dumps(v)


def decode_resource_list(rl) -> Dict[str, Q.Quantity]:
    """map[ResourceName]Quantity; any bad entry fails the whole decode."""
    out: Dict[str, Q.Quantity] = {}
    if rl is None:
        return out
    if not isinstance(rl, dict):
        raise DecodeError("resource list is not an object")
    for k, v in rl.items():
        try:
            out[k] = Q.unmarshal_json(_raw_token(v))
        except Q.QuantityError as e:
            raise DecodeError(f"bad quantity for {k}: {e}")
    return out


def parse_time(v) -> int:
    if v is None:
        return ZERO_TIME
    m = re.fullmatch(r"(\d{4})-(\d{2})-(\d{2})[Tt](\d{2}):(\d{2}):(\d{2})(\.\d+)?([Zz]|[+-]\d{2}:\d{2})", v)
    if not m:
        raise DecodeError(f"bad time {v!r}")
    y, mo, d, h, mi, s = (int(m.group(i)) for i in range(1, 7))
    t = calendar.timegm((y, mo, d, h, mi, s, 0, 0, 0))
    z = m.group(8)
    if z not in ("Z", "z"):
        sign = -1 if z[0] == "-" else 1
        t -= sign * (int(z[1:3]) * 3600 + int(z[4:6]) * 60)
    return t


# ---- decoded objects (only the fields the path reads) -----------------------------
class Container:
    def __init__(self, c: dict):
        res = c.get("resources") or {}
        self.limits = decode_resource_list(res.get("limits"))
        self.requests = decode_resource_list(res.get("requests"))


class Pod:
    def __init__(self, p: dict):
        meta = p.get("metadata") or {}
        spec = p.get("spec") or {}
        status = p.get("status") or {}
        self.name = meta.get("name") or ""
        self.namespace = meta.get("namespace") or ""
        self.labels = {k: (v or "") for k, v in (meta.get("labels") or {}).items()}
        self.owner_kinds = [(o.get("kind") or "") for o in (meta.get("ownerReferences") or [])]
        self.creation = parse_time(meta.get("creationTimestamp"))
        self.node_name = spec.get("nodeName") or ""
        self.containers = [Container(c) for c in (spec.get("containers") or [])]
        self.init_containers = [Container(c) for c in (spec.get("initContainers") or [])]   # decoded, never summed
        self.phase = status.get("phase") or ""
        self.host_ip = status.get("hostIP") or ""


class Node:
    def __init__(self, n: dict):
        meta = n.get("metadata") or {}
        status = n.get("status") or {}
        self.name = meta.get("name") or ""
        self.labels = {k: (v or "") for k, v in (meta.get("labels") or {}).items()}
        self.capacity = decode_resource_list(status.get("capacity"))
        self.allocatable = decode_resource_list(status.get("allocatable"))
        self.addresses = [((a.get("type") or ""), (a.get("address") or "")) for a in (status.get("addresses") or [])]


EMPTY_POD = Pod({})


# ---- gpu.go -------------------------------------------------------------------------
def gpu_in_container_deprecated(c: Container) -> int:
    val = c.limits.get(DEPRECATED)
    if val is None:
        return 0
    return Q.value(val)


def gpu_in_container(c: Container) -> int:
    val = c.limits.get(NVIDIA)
    if val is None:
        return gpu_in_container_deprecated(c)
    return Q.value(val)


def gpu_in_pod(pod: Pod) -> int:
    n = 0
    for c in pod.containers:
        n = wrap(n + gpu_in_container(c))
    return n


def gpu_in_active_pod(pod: Pod) -> int:
    if pod.phase == "Succeeded" or pod.phase == "Failed":
        return 0
    n = 0
    for c in pod.containers:
        n = wrap(n + gpu_in_container(c))
    return n


def gpu_in_node(node: Node) -> int:
    val = node.capacity.get(NVIDIA)
    if val is None:
        val = node.capacity.get(DEPRECATED)
        if val is None:
            return 0
    return Q.value(val)


def gpu_pods(pods: List[Pod]) -> List[Pod]:
    return [p for p in pods if gpu_in_pod(p) > 0]


# ---- top node -------------------------------------------------------------------------
def acquire_all_active_pods(podlist: dict) -> List[Pod]:
    """The field selector status.phase!=Succeeded,status.phase!=Failed, applied here
    because there is no API server."""
    pods = [Pod(p) for p in (podlist.get("items") or [])]
    return [p for p in pods if p.phase != "Succeeded" and p.phase != "Failed"]


def get_all_node_infos(nodelist: dict, all_pods: List[Pod]) -> List[Tuple[Node, List[Pod]]]:
    infos = []
    for n in (nodelist.get("items") or []):
        node = Node(n)
        pods = [p for p in all_pods if p.node_name == node.name]       # top_node.go:88-97
        infos.append((node, pods))
    return infos


def caculate_node_gpu(info) -> Tuple[int, int]:
    node, pods = info
    total = gpu_in_node(node)
    allocated = 0
    for p in pods:
        allocated = wrap(allocated + gpu_in_pod(p))
    return total, allocated


def go_int64_of_float(x: float) -> int:
    return int(x)


def is_master(node: Node) -> bool:
    return MASTER_LABEL in node.labels


RULE = "-" * 89 + "\n"


class TabWriter:
    """text/tabwriter with minwidth 0, tabwidth 0, padding 2, padchar ' ', flags 0."""

    def __init__(self, padding: int = 2):
        self.padding = padding
        self.buf = []

    def write(self, s: str) -> None:
        self.buf.append(s)

    def flush(self) -> str:
        text = "".join(self.buf)
        lines = [ln.split("\t") for ln in text.split("\n")]
        tail = lines.pop()                     # text after the last newline
        out: List[str] = []
        self._format(lines, 0, len(lines), [], out)
        out.append("".join(tail))
        return "".join(out)

    def _emit(self, lines, a, b, widths, out):
        for ln in lines[a:b]:
            for j, cell in enumerate(ln):
                out.append(cell)
                if j < len(widths):
                    out.append(" " * max(0, widths[j] - len(cell)))
            out.append("\n")

    def _format(self, lines, a, b, widths, out):
        col = len(widths)
        i = a
        start = a
        while i < b:
            if col >= len(lines[i]) - 1:       # the last cell of a line is not in a column
                i += 1
                continue
            self._emit(lines, start, i, widths, out)
            start = i
            width = 0
            while i < b and col < len(lines[i]) - 1:
                width = max(width, len(lines[i][col]) + self.padding)
                i += 1
            self._format(lines, start, i, widths + [width], out)
            start = i
        self._emit(lines, start, b, widths, out)


def display_top_node_summary(infos) -> str:
    w = TabWriter()
    total_c = alloc_c = 0
    w.write("NAME\tIPADDRESS\tROLE\tGPU(Total)\tGPU(Allocated)\n")
    for info in infos:
        node = info[0]
        total, alloc = caculate_node_gpu(info)
        total_c = wrap(total_c + total)
        alloc_c = wrap(alloc_c + alloc)
        address = "unknown"
        if len(node.addresses) > 0:
            for t, a in node.addresses:
                if t == "InternalIP":
                    address = a
                    break
        role = "master" if is_master(node) else "worker"
        w.write(f"{node.name}\t{address}\t{role}\t{total}\t{alloc}\n")
    w.write(RULE)
    w.write("Allocated/Total GPUs In Cluster:\n")
    usage = 0.0
    if total_c > 0:
        usage = float(alloc_c) / float(total_c) * 100
    w.write(f"{alloc_c}/{total_c} ({go_int64_of_float(usage)}%)\t\n")
    return w.flush()


def display_top_node_details(infos) -> str:
    w = TabWriter()
    total_c = alloc_c = 0
    w.write("\n")
    for info in infos:
        node, pods = info
        total, alloc = caculate_node_gpu(info)
        total_c = wrap(total_c + total)
        alloc_c = wrap(alloc_c + alloc)
        address = "unknown"
        if len(node.addresses) > 0:
            address = node.addresses[0][1]
        role = "master" if is_master(node) else "worker"
        w.write("\n")
        w.write(f"NAME:\t{node.name}\n")
        w.write(f"IPADDRESS:\t{address}\n")
        w.write(f"ROLE:\t{role}\n")
        gp = gpu_pods(pods)
        if len(gp) > 0:
            w.write("\n")
            w.write("NAMESPACE\tNAME\tGPU REQUESTS\tGPU LIMITS\n")
            for p in gp:
                w.write(f"{p.namespace}\t{p.name}\t{gpu_in_pod(p)}\t{gpu_in_pod(p)}\n")
            w.write("\n")
        usage = 0.0
        if total > 0:
            usage = float(alloc) / float(total) * 100
        else:
            w.write("\n")
        w.write(f"Total GPUs In Node {node.name}:\t{total} \t\n")
        w.write(f"Allocated GPUs In Node {node.name}:\t{alloc} ({go_int64_of_float(usage)}%)\t\n")
        w.write(RULE)
    w.write("\n")
    w.write("\n")
    w.write("Allocated/Total GPUs In Cluster:\t")
    usage = 0.0
    if total_c > 0:
        usage = float(alloc_c) / float(total_c) * 100
    w.write(f"{alloc_c}/{total_c} ({go_int64_of_float(usage)}%)\t\n")
    return w.flush()


def top_node(podlist_json, nodelist_json, details: bool = False):
    """Returns (text, table) where table = [(name, total, allocated)], plus totals."""
    podlist, nodelist = loads(podlist_json), loads(nodelist_json)
    infos = get_all_node_infos(nodelist, acquire_all_active_pods(podlist))
    text = display_top_node_details(infos) if details else display_top_node_summary(infos)
    table = [(i[0].name,) + caculate_node_gpu(i) for i in infos]
    return text, table


# ---- top job ----------------------------------------------------------------------------
def list_release_map(helm_text: str) -> Dict[str, str]:
    rel: Dict[str, str] = {}
    for line in helm_text.split("\n"):
        line = line.strip(" ")
        if "NAME" not in line:
            cols = line.split()
            if len(cols) > 1:
                rel[cols[0]] = cols[-1]
    return rel


class JobObj:
    def __init__(self, j: dict):
        meta = j.get("metadata") or {}
        self.name = meta.get("name") or ""
        self.namespace = meta.get("namespace") or ""
        self.labels = {k: (v or "") for k, v in (meta.get("labels") or {}).items()}
        self.status = j.get("status") or {}

    def num(self, key) -> int:
        v = self.status.get(key)
        return int(v) if v is not None else 0

    def start_time(self) -> Optional[int]:
        v = self.status.get("startTime")
        return None if v is None else parse_time(v)


def _has(labels, k, want):
    return k in labels and labels[k] == want


class TrainingJob:
    def __init__(self, name, trainer, obj, pods, chief):
        self.name, self.trainer, self.obj, self.pods, self.chief = name, trainer, obj, pods, chief

    def requested_gpu(self) -> int:
        n = 0
        for p in self.pods:
            n = wrap(n + gpu_in_pod(p))
        return n

    def allocated_gpu(self) -> int:
        n = 0
        for p in self.pods:
            n = wrap(n + gpu_in_active_pod(p))
        return n

    def get_status(self) -> str:
        if self.trainer == "tfjob":                       # trainer_tensorflow.go:69-86
            if self.obj is None or self.obj.name == "":
                return "UNKNOWN"

            def cond(t):
                return any((c.get("type") == t and c.get("status") == "True") for c in (self.obj.status.get("conditions") or []))
            if cond("Succeeded"):
                return "SUCCEEDED"
            if cond("Failed"):
                return "FAILED"
            if cond("Created") or cond("Restarting"):
                return "PENDING"
            return "RUNNING"
        status = ""                                        # job_info.go:88-108
        job = self.obj
        if job is not None:
            if job.num("active") > 0:
                status = "RUNNING"
            elif job.num("succeeded") > 0:
                status = "SUCCEEDED"
            elif job.num("failed") > 0:
                status = "FAILED"
        if status == "RUNNING":
            if self.chief.host_ip == "":
                status = "PENDING"
            elif self.chief.phase == "Pending":
                status = "PENDING"
        return status

    def host_ip_of_chief(self) -> str:
        return self.chief.host_ip if self.get_status() == "RUNNING" else "N/A"

    def age(self, now: int) -> str:
        st = self.obj.start_time() if self.obj is not None else None
        if st is None or st == ZERO_TIME:
            return "0s"
        return short_human_duration(now - st)


def short_human_duration(seconds: int) -> str:
    if seconds < -1:
        return "<invalid>"
    if seconds < 0:
        return "0s"
    if seconds < 60:
        return f"{seconds}s"
    minutes = seconds // 60
    if minutes < 60:
        return f"{minutes}m"
    hours = seconds // 3600
    if hours < 24:
        return f"{hours}h"
    if hours < 24 * 365:
        return f"{hours // 24}d"
    return f"{hours // 24 // 365}y"


def _latest_scan(cands: List[Pod]) -> Pod:
    """`if jobPod.Name == "" {first} else if latest.Before(current) {replace}`:
    strict Before, so the first pod in list order wins ties (trainer_mpi.go:263-278)."""
    job_pod = EMPTY_POD
    latest = None
    for item in cands:
        if job_pod.name == "":
            latest = item.creation
            job_pod = item
        elif latest < item.creation:
            job_pod = item
            latest = item.creation
    return job_pod


def job_list(podlist_json, joblist_json, tfjoblist_json, helm_text, now: int):
    """`arena list` (cmd/arena/commands/list.go:16-80): the same jobs as `top job`, ordered by
    makeTrainingJobOrderdByAge (trainer.go:39-66: a job without status.startTime first, then newest
    start first; ties, which sort.Sort leaves open, by name) and printed without the GPU columns and
    totals (displayTrainingJobList(jobs, false), list.go:82-133).  Returns (text, [names])."""
    return top_job(podlist_json, joblist_json, tfjoblist_json, helm_text, now, display_gpu=False)


def top_job(podlist_json, joblist_json, tfjoblist_json, helm_text, now: int, display_gpu: bool = True):
    """Returns (text, rows) with rows = [(name, trainer, requested, allocated)] in print order."""
    all_pods = [Pod(p) for p in (loads(podlist_json).get("items") or [])]            # acquireAllPods: every phase
    all_jobs = [JobObj(j) for j in (loads(joblist_json).get("items") or [])] if joblist_json else []
    tf_enabled = tfjoblist_json is not None
    all_tfjobs = [JobObj(j) for j in (loads(tfjoblist_json).get("items") or [])] if (tfjoblist_json and all_pods) else []
    if isinstance(helm_text, bytes):
        helm_text = helm_text.decode()
    releases = list_release_map(helm_text)

    def match(obj_or_pod, name, ns, app):
        return (_has(obj_or_pod.labels, "release", name) and _has(obj_or_pod.labels, "app", app)
                and obj_or_pod.namespace == ns)

    jobs: List[TrainingJob] = []
    for name, ns in releases.items():
        # MPI (trainer_mpi.go:91-119): cache path only (no API server in fixtures)
        if all_jobs and any(match(j, name, ns, "tf-horovod") for j in all_jobs):
            obj = next(j for j in all_jobs if match(j, name, ns, "tf-horovod"))
            pods, cands = [], []
            for item in all_pods:
                if not match(item, name, ns, "tf-horovod"):
                    continue
                if "Job" in item.owner_kinds:
                    cands.append(item)
                else:
                    pods.append(item)
            job_pod = _latest_scan(cands)
            pods.append(job_pod)
            jobs.append(TrainingJob(name, "mpijob", obj, pods, job_pod))
        elif all_jobs and any(match(j, name, ns, "training") for j in all_jobs):
            obj = next(j for j in all_jobs if match(j, name, ns, "training"))
            job_pod = _latest_scan([p for p in all_pods if match(p, name, ns, "training")])
            jobs.append(TrainingJob(name, "standalonejob", obj, [job_pod], job_pod))
        elif tf_enabled and all_tfjobs and any(match(j, name, ns, "tfjob") for j in all_tfjobs):
            obj = next(j for j in all_tfjobs if match(j, name, ns, "tfjob"))
            pods, chief = [], EMPTY_POD
            for item in all_pods:
                if not (match(item, name, ns, "tfjob") and _has(item.labels, "group_name", "kubeflow.org")):
                    continue
                if _has(item.labels, "tf-replica-type", "worker") and _has(item.labels, "tf-replica-index", "0"):
                    chief = item
                pods.append(item)
            jobs.append(TrainingJob(name, "tfjob", obj, pods, chief))
    if display_gpu:
        # orderedTrainingJob: requested desc; ties (unspecified in the reference) by name
        jobs.sort(key=lambda j: (-j.requested_gpu(), j.name))
    else:
        # orderedTrainingJobByAge.Less (trainer.go:45-53): nil StartTime sorts first, then After()
        jobs.sort(key=lambda j: (0, 0, j.name) if j.obj.start_time() is None else (1, -j.obj.start_time(), j.name))

    w = TabWriter()
    total_alloc = total_req = 0
    if not display_gpu:
        w.write("NAME\tSTATUS\tTRAINER\tAGE\tNODE\n")
        for j in jobs:
            w.write(f"{j.name}\t{j.get_status()}\t{j.trainer.upper()}\t{j.age(now)}\t{j.host_ip_of_chief()}\n")
        return w.flush(), [j.name for j in jobs]
    w.write("NAME\tSTATUS\tTRAINER\tAGE\tNODE\tGPU(Requests)\tGPU(Allocated)\n")
    rows = []
    for j in jobs:
        status = j.get_status()
        host_ip = j.host_ip_of_chief()
        req, alloc = j.requested_gpu(), j.allocated_gpu()
        total_alloc = wrap(total_alloc + alloc)
        total_req = wrap(total_req + req)
        w.write(f"{j.name}\t{status}\t{j.trainer.upper()}\t{j.age(now)}\t{host_ip}\t{req}\t{alloc}\n")
        rows.append((j.name, j.trainer, req, alloc))
    w.write("\n")
    w.write("\n")
    w.write("Total Allocated GPUs of Training Job:\n")
    w.write(f"{total_alloc} \t\n")
    w.write("\n")
    w.write("Total Requested GPUs of Training Job:\n")
    w.write(f"{total_req} \t\n")
    return w.flush(), rows
