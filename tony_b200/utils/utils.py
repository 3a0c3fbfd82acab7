"""Launcher utilities: container-request parsing, cluster-spec shaping, archives, shell exec.

Behavioural parity targets in the reference (tony-core/src/main/java/com/linkedin/tony/):
``util/Utils.java`` — parseKeyValue :250-270, executeShell :299-328, parseContainerRequests
:371-449, isArchive :477-495, constructTFConfig :503-520, parseClusterSpecFor{Pytorch,MXNet}
:598-637, zip/unzip :165-186; ``TonyClient.java`` buildTaskCommand :618-650;
``tensorflow/JobContainerRequest.java``.  Everything is re-expressed for a process-based local
cluster: no YARN records, no HDFS.
"""
from __future__ import annotations

import json
import os
import signal
import socket
import subprocess
import tarfile
import time
import zipfile
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterable, List, Optional

from ..conf import keys as K
from ..conf.configuration import TonyConfiguration


@dataclass
class JobContainerRequest:
    """Resource ask for one job type (reference: tensorflow/JobContainerRequest.java:10-63)."""

    job_name: str
    num_instances: int
    memory_mb: int
    vcores: int
    gpus: int
    priority: int
    node_label: Optional[str] = None
    depends_on: List[str] = field(default_factory=list)


def parse_key_value(kvs: Optional[Iterable[str]]) -> Dict[str, str]:
    """['a=b', 'c'] -> {'a': 'b', 'c': ''}; only the first '=' splits."""
    out: Dict[str, str] = {}
    for kv in kvs or []:
        kv = kv.strip()
        if not kv:
            continue
        if "=" in kv:
            k, v = kv.split("=", 1)
            out[k] = v
        else:
            out[kv] = ""
    return out


def _ensure_staged_tasks_integrity(prepare: List[str], training: List[str], all_jobs: List[str]) -> None:
    if not prepare and not training:
        return
    if not prepare:
        prepare.extend(j for j in all_jobs if j not in training)
    elif not training:
        training.extend(j for j in all_jobs if j not in prepare)
    if len(prepare) + len(training) != len(all_jobs):
        raise ValueError(
            f"cannot parse application stages: {len(prepare)} prepare-stage + {len(training)} "
            f"training-stage job types, but {len(all_jobs)} job types are configured")


def parse_container_requests(conf: TonyConfiguration, gpus_available: Optional[int] = None
                             ) -> Dict[str, JobContainerRequest]:
    """One request per job type with instances > 0; distinct priorities in job-type order;
    training-stage types depend on the tracked prepare-stage types."""
    jobs = conf.job_types()
    untracked = set(conf.untracked_job_types())
    prepare = conf.get_strings(K.PREPARE_STAGE)
    training = conf.get_strings(K.TRAINING_STAGE)
    _ensure_staged_tasks_integrity(prepare, training, jobs)
    to_depend_on = [j for j in prepare if j not in untracked]
    out: Dict[str, JobContainerRequest] = {}
    priority = 0
    for job in jobs:
        n = conf.num_instances(job)
        mem = conf.get_memory_mb(K.memory_key(job))
        vcores = conf.get_int(K.vcores_key(job), K.DEFAULT_VCORES)
        gpus = conf.get_int(K.gpus_key(job), K.DEFAULT_GPUS)
        if gpus > 0 and gpus_available is not None and gpus_available <= 0:
            raise RuntimeError(f"User requested {gpus} GPUs for job '{job}' but GPU is not available on the cluster.")
        depends = list(to_depend_on) if job in training else []
        explicit = conf.get_strings(K.depends_on_key(job))
        for d in explicit:
            if d not in depends:
                depends.append(d)
        if n > 0:
            out[job] = JobContainerRequest(job, n, mem, vcores, gpus, priority,
                                           conf.get(K.node_label_key(job)), depends)
            priority += 1
    return out


def build_task_command(python_venv: Optional[str], python_binary_path: Optional[str],
                       executes: Optional[str], task_params: Optional[str]) -> Optional[str]:
    """``[venv/]<python> <executes> <task_params>`` (truth table: TestApplicationMaster.java:13-41)."""
    if executes is None:
        return None
    cmd = executes
    if python_binary_path is not None:
        if python_binary_path.startswith("/") or python_venv is None:
            interp = python_binary_path
        else:
            interp = K.PYTHON_VENV_DIR + os.sep + python_binary_path
        cmd = interp + " " + executes
    if task_params is not None:
        cmd += " " + task_params
    return cmd


# ---- cluster-spec shaping --------------------------------------------------------------------
def construct_tf_config(cluster_spec_json: str, job_name: str, task_index: int) -> str:
    """TF_CONFIG: drop the sidecar tensorboard; drop evaluators unless we are one."""
    spec = json.loads(cluster_spec_json)
    spec.pop(K.TENSORBOARD_JOB_NAME, None)
    if job_name != K.EVALUATOR_JOB_NAME:
        spec.pop(K.EVALUATOR_JOB_NAME, None)
    return json.dumps({"cluster": spec, "task": {"type": job_name, "index": task_index}},
                      separators=(",", ":"))


def parse_cluster_spec_for_pytorch(cluster_spec_json: str) -> Optional[str]:
    spec = json.loads(cluster_spec_json)
    workers = spec.get(K.WORKER_JOB_NAME) or []
    if not workers or not workers[0]:
        return None
    return K.COMMUNICATION_BACKEND + workers[0]


def split_address_port(addr: str) -> Optional[List[str]]:
    i = addr.rfind(":")
    if i <= 0:
        return None
    return [addr[:i], addr[i + 1:]]


def parse_cluster_spec_for_mxnet(cluster_spec_json: str) -> Optional[List[str]]:
    spec = json.loads(cluster_spec_json)
    sched = spec.get(K.SCHEDULER_JOB_NAME) or []
    if not sched:
        return None
    parts = split_address_port(sched[0])
    if parts is None:
        return None
    try:
        parts[0] = socket.gethostbyname(parts[0])
    except OSError:
        return None
    return parts


# ---- archives ---------------------------------------------------------------------------------
def is_archive(path: str) -> bool:
    """Magic-number sniff for zip / gzip / tar, like the reference's isArchive."""
    try:
        with open(path, "rb") as f:
            head = f.read(512)
    except OSError:
        return False
    if head[:4] in (b"PK\x03\x04", b"PK\x05\x06", b"PK\x07\x08"):
        return True
    if head[:2] == b"\x1f\x8b":
        return True
    return len(head) >= 262 and head[257:262] == b"ustar"


def zip_folder(src_dir: str, dst_zip: str) -> None:
    src_dir = os.path.abspath(src_dir)
    with zipfile.ZipFile(dst_zip, "w", zipfile.ZIP_DEFLATED) as z:
        for root, _, files in os.walk(src_dir):
            for name in files:
                full = os.path.join(root, name)
                z.write(full, os.path.relpath(full, src_dir))


def unzip_archive(src: str, dst_dir: str) -> None:
    os.makedirs(dst_dir, exist_ok=True)
    if zipfile.is_zipfile(src):
        with zipfile.ZipFile(src) as z:
            for info in z.infolist():
                z.extract(info, dst_dir)
                mode = info.external_attr >> 16
                if mode:
                    try:
                        os.chmod(os.path.join(dst_dir, info.filename), mode)
                    except OSError:
                        pass
    elif tarfile.is_tarfile(src):
        with tarfile.open(src) as t:
            t.extractall(dst_dir)
    else:
        raise ValueError(f"{src} is not a zip or tar archive")


# ---- polling / processes ----------------------------------------------------------------------
def poll(fn: Callable[[], bool], interval_s: float, timeout_s: float) -> bool:
    """Call ``fn`` every ``interval_s`` until it returns True or ``timeout_s`` elapsed (<=0: forever)."""
    deadline = None if timeout_s <= 0 else time.monotonic() + timeout_s
    while True:
        if fn():
            return True
        if deadline is not None and time.monotonic() >= deadline:
            return False
        time.sleep(interval_s)


def poll_till_non_null(fn: Callable[[], Optional[object]], interval_s: float, timeout_s: float):
    deadline = None if timeout_s <= 0 else time.monotonic() + timeout_s
    while True:
        r = fn()
        if r is not None:
            return r
        if deadline is not None and time.monotonic() >= deadline:
            return None
        time.sleep(interval_s)


def execute_shell(command: str, timeout_ms: int = 0, env: Optional[Dict[str, str]] = None,
                  cwd: Optional[str] = None, stdout=None, stderr=None,
                  on_start: Optional[Callable[[subprocess.Popen], None]] = None) -> int:
    """``bash -c command`` with inherited IO; MALLOC_ARENA_MAX dropped; whole process group
    killed on timeout (exit code -1).  Reference: util/Utils.java:299-328."""
    full_env = dict(os.environ)
    full_env.pop("MALLOC_ARENA_MAX", None)
    if env:
        full_env.update(env)
    proc = subprocess.Popen(["bash", "-c", command], env=full_env, cwd=cwd, stdout=stdout,
                            stderr=stderr, start_new_session=True)
    if on_start:
        on_start(proc)
    try:
        return proc.wait(timeout=timeout_ms / 1000.0 if timeout_ms and timeout_ms > 0 else None)
    except subprocess.TimeoutExpired:
        kill_process_group(proc)
        return -1


def kill_process_group(proc: subprocess.Popen, grace_s: float = 2.0) -> None:
    """SIGTERM then SIGKILL the exact process group we started (never by pattern)."""
    try:
        pgid = os.getpgid(proc.pid)
    except ProcessLookupError:
        return
    for sig, wait in ((signal.SIGTERM, grace_s), (signal.SIGKILL, 1.0)):
        try:
            os.killpg(pgid, sig)
        except ProcessLookupError:
            return
        try:
            proc.wait(timeout=wait)
            return
        except subprocess.TimeoutExpired:
            continue


def local_hostname() -> str:
    """The address tasks on this box use to reach each other (single node: loopback-safe)."""
    return os.environ.get("TONY_B200_HOSTNAME", "127.0.0.1")


def task_id(job_name: str, index: int) -> str:
    return f"{job_name}:{index}"


def current_user() -> str:
    import getpass

    try:
        u = getpass.getuser()
    except Exception:
        u = "tony"
    # history file names forbid '-' and '.' in user names and require lower case
    return "".join(c for c in u.lower() if c.isalnum() or c == "_") or "tony"
