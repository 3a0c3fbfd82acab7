"""Job-type DAG scheduler: release container requests as their dependencies complete.

Reference: tony-core/src/main/java/com/linkedin/tony/TaskScheduler.java — dependency map from
``dependsOn`` :75-86, cycle check :142-178, schedule ready types :55-73,93-105, release
dependents when every instance of a type has completed :118-140.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Set

from ..utils.utils import JobContainerRequest
from .session import TonySession


def is_dag(requests: Dict[str, JobContainerRequest]) -> bool:
    """True when the depends-on graph over job types has no cycle (and no unknown type)."""
    WHITE, GREY, BLACK = 0, 1, 2
    color = {j: WHITE for j in requests}

    def visit(j: str) -> bool:
        color[j] = GREY
        for d in requests[j].depends_on:
            if d not in requests:
                continue  # dependency on a type with zero instances is vacuous
            if color[d] == GREY:
                return False
            if color[d] == WHITE and not visit(d):
                return False
        color[j] = BLACK
        return True

    return all(color[j] != WHITE or visit(j) for j in requests)


class TaskScheduler:
    def __init__(self, session: TonySession, request_containers: Callable[[JobContainerRequest], None]):
        self.session = session
        self.request_containers = request_containers
        self.requests = session.requests
        self.dependency_check_passed = True
        # job -> set of job types it still waits for
        self.waiting: Dict[str, Set[str]] = {}
        self.remaining_instances: Dict[str, int] = {j: r.num_instances for j, r in self.requests.items()}
        self.scheduled: List[str] = []

    def schedule_tasks(self) -> None:
        if not is_dag(self.requests):
            self.dependency_check_passed = False
            return
        for job, req in self.requests.items():
            deps = {d for d in req.depends_on if d in self.requests}
            if deps:
                self.waiting[job] = deps
        for job, req in self.requests.items():
            if job not in self.waiting:
                self._schedule(req)

    def _schedule(self, req: JobContainerRequest) -> None:
        self.scheduled.append(req.job_name)
        # the gang barrier only counts job types the DAG has released so far
        self.session.num_expected_tasks += req.num_instances
        self.request_containers(req)

    def register_dependency_completed(self, job: str) -> None:
        """One instance of ``job`` finished; when all did, release the types waiting on it."""
        if job not in self.remaining_instances:
            return
        self.remaining_instances[job] -= 1
        if self.remaining_instances[job] > 0:
            return
        for waiter in list(self.waiting):
            deps = self.waiting[waiter]
            deps.discard(job)
            if not deps:
                del self.waiting[waiter]
                self._schedule(self.requests[waiter])
