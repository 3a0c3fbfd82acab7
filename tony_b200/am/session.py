"""In-memory state of one application attempt: tasks, cluster spec, completion policy.

Policy parity with the reference (tony-core/src/main/java/com/linkedin/tony/tensorflow/
TonySession.java): chief = ``chief:*`` or ``worker:0`` when no chief type :383-386; a non-zero
exit (other than killed-by-AM) of the chief, of a stop-on-failure type, or of anything when
fail-on-worker-failure is on, ends training as FAILED at once :262-286; otherwise the final
status is FAILED only if every tracked task failed :295-349; statuses per exit code :506-523.
Unlike the reference (which mutates plain HashSets from RPC threads) all
state here is guarded by the ApplicationMaster's single lock.
"""
from __future__ import annotations

import json
import time
from dataclasses import dataclass, field
from enum import IntEnum
from typing import Dict, List, Optional, Set

from ..conf import keys as K
from ..conf.configuration import TonyConfiguration
from ..utils.utils import JobContainerRequest, task_id


class TaskStatus(IntEnum):
    """Wire-significant ordinals (yarn_tensorflow_cluster_protos.proto:16-23)."""

    NEW = 0
    READY = 1
    RUNNING = 2
    FAILED = 3
    SUCCEEDED = 4
    FINISHED = 5


# attention order for the client's task table: failures first (rpc/TaskInfo.java:48-58)
_ATTENTION = {TaskStatus.FAILED: 0, TaskStatus.SUCCEEDED: 1, TaskStatus.FINISHED: 2,
              TaskStatus.RUNNING: 3, TaskStatus.NEW: 4, TaskStatus.READY: 5}


class FinalStatus:
    UNDEFINED = "UNDEFINED"
    SUCCEEDED = "SUCCEEDED"
    FAILED = "FAILED"
    KILLED = "KILLED"


@dataclass(order=False)
class TaskInfo:
    """Client-visible task row."""

    name: str
    index: int
    url: str = ""
    status: TaskStatus = TaskStatus.NEW

    def sort_key(self):
        return (_ATTENTION[self.status], self.name, self.index)

    def to_dict(self):
        return {"name": self.name, "index": self.index, "url": self.url, "status": int(self.status)}

    @staticmethod
    def from_dict(d):
        return TaskInfo(d["name"], int(d["index"]), d.get("url", ""), TaskStatus(int(d["status"])))

    def __hash__(self):
        return hash((self.name, self.index))

    def __eq__(self, other):
        return isinstance(other, TaskInfo) and (self.name, self.index) == (other.name, other.index)


@dataclass
class TonyTask:
    job_name: str
    index: int
    session_id: int
    start_time_ms: int = field(default_factory=lambda: int(time.time() * 1000))
    host: Optional[str] = None
    port: int = -1
    container_id: Optional[str] = None
    url: str = ""
    status: TaskStatus = TaskStatus.NEW
    exit_status: Optional[int] = None
    completed: bool = False
    metrics: Dict[str, float] = field(default_factory=dict)

    @property
    def id(self) -> str:
        return task_id(self.job_name, self.index)

    @property
    def host_port(self) -> Optional[str]:
        return None if self.host is None else f"{self.host}:{self.port}"

    def set_host_port(self, spec: str) -> None:
        host, _, port = spec.rpartition(":")
        self.host, self.port = host, int(port)

    def set_exit_status(self, code: int) -> None:
        """First terminal state wins (a late container callback must not overwrite it)."""
        if self.completed:
            return
        self.exit_status = code
        self.completed = True
        if code == 0:
            self.status = TaskStatus.SUCCEEDED
        elif code == K.KILLED_BY_APPMASTER:
            self.status = TaskStatus.FINISHED
        else:
            self.status = TaskStatus.FAILED

    @property
    def failed(self) -> bool:
        return self.status == TaskStatus.FAILED

    def info(self) -> TaskInfo:
        return TaskInfo(self.job_name, self.index, self.url, self.status)


class TonySession:
    def __init__(self, conf: TonyConfiguration, requests: Dict[str, JobContainerRequest], session_id: int = 0):
        self.conf = conf
        self.session_id = session_id
        self.requests = requests
        self.job_tasks: Dict[str, List[Optional[TonyTask]]] = {
            job: [None] * req.num_instances for job, req in requests.items()}
        self.registered: Set[str] = set()
        self.num_expected_tasks = 0  # only job types the DAG scheduler has released so far
        self.training_finished = False
        self.final_status = FinalStatus.UNDEFINED
        self.final_message: Optional[str] = None
        self.untracked = set(conf.untracked_job_types())
        self.sidecars = set(conf.sidecar_job_types())
        self.stop_on_failure = set(conf.get_strings(K.STOP_ON_FAILURE_JOBTYPES))
        self.fail_on_worker_failure = conf.get_bool(K.FAIL_ON_WORKER_FAILURE, False)

    # -- construction of tasks ------------------------------------------------------------
    def init_task_by_priority(self, priority: int) -> Optional[TonyTask]:
        """Bind the next unassigned slot of the job type with this priority."""
        for job, req in self.requests.items():
            if req.priority != priority:
                continue
            slots = self.job_tasks[job]
            for i, t in enumerate(slots):
                if t is None:
                    slots[i] = TonyTask(job, i, self.session_id)
                    return slots[i]
        return None

    def get_task(self, job: str, index) -> Optional[TonyTask]:
        slots = self.job_tasks.get(job)
        i = int(index)
        if slots is None or i < 0 or i >= len(slots):
            return None
        return slots[i]

    def get_task_by_id(self, tid: str) -> Optional[TonyTask]:
        job, _, idx = tid.partition(":")
        return self.get_task(job, idx) if idx.isdigit() else None

    def get_task_by_container(self, container_id: str) -> Optional[TonyTask]:
        for t in self.all_tasks():
            if t.container_id == container_id:
                return t
        return None

    def all_tasks(self) -> List[TonyTask]:
        return [t for slots in self.job_tasks.values() for t in slots if t is not None]

    # -- counters -------------------------------------------------------------------------
    def is_tracked(self, job: str) -> bool:
        return job not in self.untracked and job not in self.sidecars

    def all_tasks_scheduled(self) -> bool:
        return all(t is not None for slots in self.job_tasks.values() for t in slots)

    def total_tasks(self) -> int:
        return sum(len(s) for s in self.job_tasks.values())

    def total_tracked_tasks(self) -> int:
        return sum(len(s) for j, s in self.job_tasks.items() if self.is_tracked(j))

    def num_completed_tasks(self) -> int:
        return sum(1 for t in self.all_tasks() if t.completed)

    def num_completed_tracked_tasks(self) -> int:
        return sum(1 for t in self.all_tasks() if t.completed and self.is_tracked(t.job_name))

    def num_failed_tasks(self) -> int:
        return sum(1 for t in self.all_tasks() if t.failed)

    def num_registered(self) -> int:
        return len(self.registered)

    # -- cluster spec -----------------------------------------------------------------------
    def cluster_spec(self) -> Dict[str, List[str]]:
        return {job: [t.host_port for t in slots if t is not None and t.host_port is not None]
                for job, slots in self.job_tasks.items()}

    def cluster_spec_json(self) -> str:
        return json.dumps(self.cluster_spec(), separators=(",", ":"))

    def task_infos(self) -> List[TaskInfo]:
        """Empty until every task has a container (ApplicationMaster.java:880-886)."""
        if not self.all_tasks_scheduled():
            return []
        return [t.info() for t in self.all_tasks()]

    # -- policy ---------------------------------------------------------------------------
    def is_chief(self, job: str, index) -> bool:
        return job == K.CHIEF_JOB_NAME or (
            K.CHIEF_JOB_NAME not in self.job_tasks and job == K.WORKER_JOB_NAME and int(index) == 0)

    def set_final_status(self, status: str, message: Optional[str]) -> None:
        self.final_status = status
        self.final_message = message

    def on_task_completed(self, job: str, index, exit_code: int, diagnostics: Optional[str] = None) -> None:
        task = self.get_task(job, index)
        if task is None:
            return
        task.set_exit_status(exit_code)
        if exit_code not in (0, K.KILLED_BY_APPMASTER):
            if self.is_chief(job, index) or job in self.stop_on_failure or self.fail_on_worker_failure:
                self.training_finished = True
                msg = f"Exit status: {exit_code}"
                if diagnostics:
                    msg += f". Error msg: {diagnostics}"
                self.set_final_status(FinalStatus.FAILED, msg)

    def update_session_status(self) -> None:
        if self.final_status == FinalStatus.FAILED:
            return
        failures = 0
        for job, slots in self.job_tasks.items():
            if not self.is_tracked(job):
                continue
            for t in slots:
                if t is None:
                    self.set_final_status(FinalStatus.FAILED, "Job is null, this should not happen.")
                    return
                if not t.completed:
                    self.set_final_status(FinalStatus.FAILED, f"Job {t.id} hasn't finished yet.")
                    return
                if t.exit_status != 0:
                    failures += 1
        if failures > 0:
            if self.fail_on_worker_failure or failures >= self.total_tracked_tasks():
                self.set_final_status(
                    FinalStatus.FAILED,
                    f"At least one job task exited with non-zero status, failedCnt={failures}")
            else:
                self.set_final_status(
                    FinalStatus.SUCCEEDED,
                    f"Training completed with some worker jobs failure, failedCnt={failures}")
        else:
            self.set_final_status(FinalStatus.SUCCEEDED, None)
