"""ApplicationMaster: the per-application driver process.

Lifecycle parity with tony-core/src/main/java/com/linkedin/tony/ApplicationMaster.java:
init :229-314 -> prepare (RPC servers, liveliness monitor, history) :443-527 ->
``do { start(); monitor(); if ok or no retries left: break; reset(); } while (!singleNode)``
:382-422 -> stop :735-777.  Container launch env/command :1144-1223, completion handling
:1234-1269, heartbeat expiry :1225-1232, registration timeout + startup failure :1271-1329,
preprocessing / single-node mode :580-601,780-832, RPC surface :854-970.

Re-designed, not translated: one process-wide lock + condition variable guards all state
(the reference mutates unsynchronised collections from RPC threads); the
monitor sleeps on that condition and wakes the instant anything changes (no 5 s poll);
``registerWorkerSpec`` is a long-poll released the moment the gang barrier opens (no 3 s
executor re-poll); containers are local process groups placed one per GPU.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import threading
import time
import traceback
from typing import Dict, List, Optional

from ..cluster import topology
from ..cluster.local_rm import Container, LocalResourceManager
from ..conf import keys as K
from ..conf.configuration import TonyConfiguration
from ..events import schema as ev
from ..events.history import EventHandler
from ..rpc.server import APPLICATION_RPC_METHODS, METRICS_RPC_METHODS, RpcServer
from ..runtimes.frameworks import get_am_adapter
from ..utils import docker as docker_util
from ..utils import utils as U
from ..utils.localizable_resource import parse_resources
from .liveliness import LivelinessMonitor
from .scheduler import TaskScheduler
from .session import FinalStatus, TaskStatus, TonySession

LONG_POLL_S = 20.0
REPO_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def log(msg: str) -> None:
    sys.stdout.write(f"{time.strftime('%Y-%m-%d %H:%M:%S')} AM: {msg}\n")
    sys.stdout.flush()


class ApplicationMaster:
    def __init__(self, app_dir: str, app_id: str, user: Optional[str] = None):
        self.app_dir = os.path.abspath(app_dir)
        self.app_id = app_id
        self.user = user or U.current_user()
        self.conf = TonyConfiguration.from_final_xml(os.path.join(self.app_dir, K.TONY_FINAL_XML))
        self.lock = threading.Condition(threading.RLock())
        self.session: Optional[TonySession] = None
        self.scheduler: Optional[TaskScheduler] = None
        self.session_id = 0
        self.num_am_retries = 0
        self.container_env: Dict[str, str] = {}
        self.metrics: Dict[str, Dict[str, float]] = {}
        self.first_step: Dict[str, float] = {}
        self.client_signal_to_stop = False
        self.task_has_missed_hb = False
        self.untracked_task_failed = False
        self.preprocess_exit_code = 0
        self.preprocess_finished = False
        self.single_node = False
        self.tb_url: Optional[str] = None
        self.app_state = "RUNNING"
        self.final_status = FinalStatus.UNDEFINED
        self.final_message: Optional[str] = None
        self.started_ms = int(time.time() * 1000)
        self.submit_time = float(os.environ.get("TONY_B200_SUBMIT_TIME", "0") or 0) or time.time()
        self.host = U.local_hostname()
        self.timeline: List[dict] = []   # (event, seconds since submit) for submit-to-first-step analysis
        self.rm: Optional[LocalResourceManager] = None
        self.hb: Optional[LivelinessMonitor] = None
        self.events: Optional[EventHandler] = None

    def mark(self, event: str, **extra) -> None:
        """Launcher timeline: monotonic milestones relative to the client's submit time."""
        row = {"event": event, "t_ms": round(1000.0 * (time.time() - self.submit_time), 1)}
        row.update(extra)
        self.timeline.append(row)

    # ================================================================== init / prepare
    def init(self) -> bool:
        conf = self.conf
        self.framework = conf.get(K.FRAMEWORK, "tensorflow").lower()
        self.adapter = get_am_adapter(self.framework)
        if not self.adapter.validate_and_update_config(conf):
            self.final_message = self.adapter.app_failed_reason
            log(f"config rejected by the {self.framework} runtime: {self.final_message}")
            return False
        self.distributed_mode = conf.get(K.DISTRIBUTED_MODE, "GANG").upper()
        self.app_timeout_ms = conf.get_int(K.APPLICATION_TIMEOUT, 0)
        self.registration_timeout_ms = conf.get_int(K.CONTAINER_ALLOCATION_TIMEOUT, -1)
        self.max_retries = conf.get_int(K.AM_RETRY_COUNT, 0)
        self.hb_interval_ms = conf.get_int(K.HEARTBEAT_INTERVAL_MS, 1000)
        self.max_missed_hb = conf.get_int(K.MAX_MISSED_HEARTBEATS, 25)
        self.enable_preprocess = conf.get_bool(K.ENABLE_PREPROCESS, False)
        self.single_node = conf.total_instances() == 0
        self.security = conf.get_bool(K.SECURITY_ENABLED, True)
        secret_hex = os.environ.get(K.TONY_SECRET)
        self.secret = bytes.fromhex(secret_hex) if (self.security and secret_hex) else None
        for k, v in U.parse_key_value(conf.get_strings(K.CONTAINERS_ENVS)).items():
            self.container_env[k] = v
        return True

    def prepare(self) -> None:
        self.mark("am_started")
        self.app_rpc = RpcServer(self, APPLICATION_RPC_METHODS, self.secret).start()
        self.metrics_rpc = RpcServer(self, METRICS_RPC_METHODS, self.secret).start()
        self.container_env[K.AM_HOST] = self.host
        self.container_env[K.AM_PORT] = str(self.app_rpc.port)
        self.container_env[K.METRICS_RPC_PORT] = str(self.metrics_rpc.port)
        self.hb = LivelinessMonitor(self.hb_interval_ms, self.max_missed_hb, self._on_task_deemed_dead)
        self.hb.start()
        work_root = os.path.dirname(self.app_dir)
        self.rm = LocalResourceManager(self.conf, self.app_id, os.path.join(work_root, "cluster"),
                                       self._on_container_allocated, self._on_container_completed)
        # history: job dir holds tony-final.xml + the jhist (ApplicationMaster.java:535-569)
        hist_root = self.conf.get(K.HISTORY_INTERMEDIATE)
        if not hist_root or hist_root.startswith("/path/to/"):
            hist_root = os.path.join(work_root, "history", "intermediate")
        self.job_dir = os.path.join(hist_root, self.app_id)
        os.makedirs(self.job_dir, exist_ok=True)
        self.conf.write_xml(os.path.join(self.job_dir, K.TONY_FINAL_XML))
        self.events = EventHandler(self.job_dir, self.app_id, self.started_ms, self.user)
        self.events.start()
        try:
            self.topology_json = topology.topology_json()
        except Exception:  # noqa: BLE001
            self.topology_json = "{}"
        with open(os.path.join(self.app_dir, "am.json.tmp"), "w") as f:
            json.dump({"host": self.host, "port": self.app_rpc.port, "metrics_port": self.metrics_rpc.port,
                       "pid": os.getpid(), "job_dir": self.job_dir}, f)
        os.replace(os.path.join(self.app_dir, "am.json.tmp"), os.path.join(self.app_dir, "am.json"))

    # ================================================================== RPC surface
    def getTaskInfos(self):
        with self.lock:
            if self.single_node and self.tb_url:
                # notebook / single-node mode: surface the AM-hosted service as a task row
                return [{"name": K.NOTEBOOK_JOB_NAME, "index": 0, "url": self.tb_url,
                         "status": int(TaskStatus.RUNNING)}]
            if self.session is None:
                return []
            return [t.to_dict() for t in self.session.task_infos()]

    def getClusterSpec(self):
        with self.lock:
            return self.session.cluster_spec_json() if self.session else "{}"

    def registerWorkerSpec(self, worker: str, spec: str):
        deadline = time.monotonic() + LONG_POLL_S
        with self.lock:
            session = self.session
            if session is None:
                return None
            task = session.get_task_by_id(worker)
            if task is None:
                return None
            if task.host is None:
                task.set_host_port(spec)
                session.registered.add(worker)
                self.hb.register(worker)
                log(f"registered {worker} at {spec} ({session.num_registered()}/{session.num_expected_tasks})")
                self.mark("task_registered", task=worker)
                if session.num_registered() == session.num_expected_tasks:
                    self.mark("gang_barrier_open")
                self._kill_chief_worker_if_testing(worker)
                self.lock.notify_all()
            while True:
                if self.session is not session or session.training_finished or self.app_state != "RUNNING":
                    return None
                if self.adapter.can_start_task(self.distributed_mode, worker):
                    out = self.adapter.construct_cluster_spec(worker)
                    if out is not None:
                        return out
                elif self.adapter.app_failed_reason:
                    session.set_final_status(FinalStatus.FAILED, self.adapter.app_failed_reason)
                    session.training_finished = True
                    self.lock.notify_all()
                    return None
                remaining = deadline - time.monotonic()
                if remaining <= 0:
                    return None
                self.lock.wait(remaining)

    def registerTensorBoardUrl(self, spec: str):
        with self.lock:
            self.tb_url = spec if spec.startswith("http") else f"http://{spec}"
            log(f"tracking URL: {self.tb_url}")
            self.lock.notify_all()
        return "SUCCEEDED"

    def registerExecutionResult(self, exitCode: int, jobName: str, jobIndex: str, sessionId: str):
        # Unregister from the liveliness monitor first so a late container-exit callback
        # cannot race with heartbeat expiry (ApplicationMaster.java:928-956).
        with self.lock:
            if int(sessionId) == self.session_id:
                self.hb.unregister(U.task_id(jobName, int(jobIndex)))
        return "RECEIVED"

    def finishApplication(self):
        with self.lock:
            self.client_signal_to_stop = True
            self.lock.notify_all()
        return None

    def taskExecutorHeartbeat(self, taskId: str):
        self.hb.received_ping(taskId)
        return None

    def registerCallbackInfo(self, taskId: str, callbackInfo: str):
        with self.lock:
            ok = self.adapter.receive_task_callback_info(taskId, callbackInfo)
            self.lock.notify_all()
        return ok

    def updateMetrics(self, taskType: str, taskIndex: int, metrics: List[dict]):
        with self.lock:
            tid = U.task_id(taskType, int(taskIndex))
            cur = self.metrics.setdefault(tid, {})
            cur.update({m["name"]: float(m["value"]) for m in metrics})
        return None

    def markFirstStep(self, taskId: str, timestamp: float):
        with self.lock:
            if taskId not in self.first_step:
                self.first_step[taskId] = float(timestamp)
                self.timeline.append({"event": "first_step", "task": taskId,
                                      "t_ms": round(1000.0 * (timestamp - self.submit_time), 1)})
                log(f"first step of {taskId}: {1000.0 * (timestamp - self.submit_time):.0f} ms after submit")
                # lands in the task's TaskFinished.metrics next to the resource gauges
                self.metrics.setdefault(taskId, {})["SUBMIT_TO_FIRST_STEP_MS"] = \
                    round(1000.0 * (float(timestamp) - self.submit_time), 1)
        return None

    def getApplicationStatus(self):
        with self.lock:
            latency = None
            if self.first_step:
                latency = 1000.0 * (min(self.first_step.values()) - self.submit_time)
            return {"state": self.app_state, "finalStatus": self.final_status, "message": self.final_message,
                    "sessionId": self.session_id, "trackingUrl": self.tb_url,
                    "submitToFirstStepMs": latency, "jobDir": getattr(self, "job_dir", None)}

    # ================================================================== containers
    def _task_env(self, task, container: Container) -> Dict[str, str]:
        s = self.session
        env = dict(self.container_env)
        env[K.JOB_NAME] = task.job_name
        env[K.JOB_ID] = self.app_id
        env[K.APP_ID] = self.app_id
        env[K.TASK_INDEX] = str(task.index)
        env[K.TASK_NUM] = str(s.total_tracked_tasks())
        env[K.DISTRIBUTED_MODE_NAME] = self.distributed_mode
        if s.is_chief(task.job_name, task.index):
            env[K.IS_CHIEF] = "true"
        env[K.SESSION_ID] = str(self.session_id)
        env[K.ATTEMPT_NUMBER] = "0"
        if self.num_am_retries:
            env[K.NUM_AM_RETRIES] = str(self.num_am_retries)
        env[K.CONTAINER_ID] = container.container_id
        env[K.TONY_CONF_PATH] = os.path.join(self.app_dir, K.TONY_FINAL_XML)
        if self.secret:
            env[K.TONY_SECRET] = self.secret.hex()
        env[K.TONY_TOPOLOGY] = self.topology_json
        env["PYTHONPATH"] = REPO_ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")
        env["PYTHONUNBUFFERED"] = "1"
        env.update(docker_util.container_runtime_env(self.conf, task.job_name))
        return env

    def _localize(self, job_name: str, container_dir: str) -> None:
        os.makedirs(container_dir, exist_ok=True)
        for key in (K.CONTAINERS_RESOURCES, K.resources_key(job_name)):
            for r in parse_resources(self.conf.get(key)):
                r.localize(container_dir)

    def _on_container_allocated(self, container: Container) -> None:
        with self.lock:
            session = self.session
            task = session.init_task_by_priority(container.priority) if session else None
            if task is None:
                log(f"no task wants {container.container_id}; releasing it")
                self.rm.release_unlaunched(container)
                return
            task.container_id = container.container_id
            task.status = TaskStatus.READY
            container_dir = os.path.join(self.app_dir, "containers", container.container_id)
            task.url = f"file://{container_dir}/stdout"
            env = self._task_env(task, container)
        try:
            self._localize(task.job_name, container_dir)
            cmd = [sys.executable, "-m", "tony_b200.executor.task_executor"]
            docker_util.write_env_file(self.conf, env, container_dir)
            cmd = docker_util.wrap_command(self.conf, task.job_name, cmd, container, container_dir)
            self.rm.start_container(container, cmd, env, container_dir)
        except Exception as exc:  # noqa: BLE001
            log(f"failed to launch {task.id}: {exc}")
            with self.lock:
                session.on_task_completed(task.job_name, task.index, 1, f"launch failed: {exc}")
                self.rm.release_unlaunched(container)
                self.lock.notify_all()
            return
        with self.lock:
            task.status = TaskStatus.RUNNING
            log(f"launched {task.id} in {container.container_id} gpus={container.gpu_ids}")
            self.mark("task_launched", task=task.id)
            self.events.emit(ev.task_started(task.job_name, task.index, container.host, container.container_id))
            self.lock.notify_all()

    def _on_container_completed(self, container: Container, exit_status: int, diagnostics: str) -> None:
        if os.environ.get(K.TEST_TASK_COMPLETION_NOTIFICATION_DELAYED):
            time.sleep(1.0)  # fault hook: late completion callback (ApplicationMaster.java:1074-1083)
        with self.lock:
            session = self.session
            task = session.get_task_by_container(container.container_id) if session else None
            if task is None:
                return  # stale completion from a previous session
            self.hb.unregister(task.id)
            session.on_task_completed(task.job_name, task.index, exit_status, diagnostics)
            if self.scheduler:
                self.scheduler.register_dependency_completed(task.job_name)
            log(f"{task.id} completed with exit status {exit_status} -> {task.status.name}")
            if exit_status != 0:
                self.mark("task_failed", task=task.id, exit=exit_status)
            self.events.emit(ev.task_finished(task.job_name, task.index, task.status.name,
                                              self.metrics.get(task.id), diagnostics or None))
            if task.job_name in session.untracked and task.failed:
                self.untracked_task_failed = True
            self.lock.notify_all()

    def _on_task_deemed_dead(self, task_id: str) -> None:
        with self.lock:
            log(f"task {task_id} missed too many heartbeats")
            task = self.session.get_task_by_id(task_id) if self.session else None
            if task is not None and not task.completed:
                self.task_has_missed_hb = True
                task.status = TaskStatus.FAILED
            self.lock.notify_all()

    def _kill_chief_worker_if_testing(self, task_id: str) -> None:
        """Fault hook TEST_WORKER_TERMINATION: crash the workers once worker:0 has registered
        (first attempt only, so a retry can succeed)."""
        if not os.environ.get(K.TEST_WORKER_TERMINATED) or task_id != "worker:0" or self.num_am_retries:
            return
        for t in self.session.all_tasks():
            if t.job_name == K.WORKER_JOB_NAME and t.container_id:
                c = self.rm.container(t.container_id)
                if c is not None and c.proc is not None and c.proc.poll() is None:
                    log(f"simulating termination of {t.id}")
                    try:
                        os.killpg(c.proc.pid, 9)
                    except ProcessLookupError:
                        pass

    # ================================================================== session driving
    def _build_session(self) -> Optional[str]:
        reqs = U.parse_container_requests(self.conf, gpus_available=len(self.rm.inv.gpus))
        for r in reqs.values():
            err = self.rm.validate(r)
            if err:
                return err
        self.session = TonySession(self.conf, reqs, self.session_id)
        self.adapter.set_session(self.session)
        self.scheduler = TaskScheduler(self.session, self.rm.add_container_request)
        return None

    def _run_am_command(self) -> int:
        """Preprocessing / single-node execution inside the AM (ApplicationMaster.java:780-832)."""
        cmd = self.conf.get(K.AM_COMMAND) or self.conf.get(K.CONTAINERS_COMMAND)
        if not cmd:
            return 0
        env = U.parse_key_value(self.conf.get_strings(K.EXECUTION_ENVS))
        env[K.PREPROCESSING_JOB] = "true"
        env["HOME"] = self.app_dir
        env["PYTHONPATH"] = REPO_ROOT + os.pathsep + os.environ.get("PYTHONPATH", "")
        tb = None
        if self.single_node:
            from ..executor.ports import EphemeralPort

            tb = EphemeralPort.create()
            env[K.TB_PORT] = str(tb.port)
            self.registerTensorBoardUrl(f"{self.host}:{tb.port}")
            tb.close()
        out_path = os.path.join(self.app_dir, "am_command.stdout")
        with open(out_path, "ab") as out:
            code = U.execute_shell(cmd, self.app_timeout_ms, env=env, cwd=self.app_dir, stdout=out,
                                   stderr=subprocess.STDOUT)
        try:
            with open(out_path, errors="replace") as f:
                for line in f:
                    m = re.search(r"Model parameters: (.+)", line)
                    if m:
                        self.container_env[K.MODEL_PARAMS] = m.group(1).strip()
        except OSError:
            pass
        return code

    def start(self) -> bool:
        if self.enable_preprocess or self.single_node:
            self.preprocess_exit_code = self._run_am_command()
            self.preprocess_finished = True
            if self.preprocess_exit_code != 0 or self.single_node:
                if self.session is None:
                    self.session = TonySession(self.conf, {}, self.session_id)
                return self.preprocess_exit_code == 0
        with self.lock:
            err = self._build_session()
            if err:
                self.session = TonySession(self.conf, {}, self.session_id)
                self.session.set_final_status(FinalStatus.FAILED, err)
                self.session.training_finished = True
                return False
            self.scheduler.schedule_tasks()
            self.lock.notify_all()
        return True

    def _registration_timed_out(self) -> bool:
        if self.registration_timeout_ms <= 0:
            return False
        now = int(time.time() * 1000)
        for t in self.session.all_tasks():
            if t.host is None and not t.completed and now - t.start_time_ms > self.registration_timeout_ms:
                self.session.set_final_status(
                    FinalStatus.FAILED,
                    f"Stopping AM for task [{t.id}] registration timeout: allocated container is {t.container_id}")
                return True
        return False

    def _startup_failed(self) -> bool:
        for t in self.session.all_tasks():
            if t.failed and t.id not in self.session.registered:
                self.session.set_final_status(
                    FinalStatus.FAILED,
                    f"Stopping AM for task [{t.id}] starting failed, allocated container is {t.container_id}")
                return True
        return False

    def monitor(self) -> bool:
        deadline = None if self.app_timeout_ms <= 0 else time.monotonic() + self.app_timeout_ms / 1000.0
        s = self.session
        with self.lock:
            while True:
                if deadline is not None and time.monotonic() > deadline:
                    log("Application times out.")
                    s.set_final_status(FinalStatus.FAILED, "Application timed out.")
                    break
                if self.client_signal_to_stop:
                    log("Client signals AM to exit.")
                    break
                if s.training_finished:
                    log("Training has finished.")
                    break
                if self.preprocess_exit_code != 0:
                    s.set_final_status(FinalStatus.FAILED, f"Preprocess failed with exit code: {self.preprocess_exit_code}")
                    break
                if self.single_node and self.preprocess_finished:
                    break
                if self.task_has_missed_hb:
                    s.set_final_status(FinalStatus.FAILED, "Application failed due to missed heartbeats")
                    break
                if self.untracked_task_failed:
                    s.set_final_status(FinalStatus.FAILED, "One of the untracked tasks has failed with a non-zero exit code.")
                    break
                if self.scheduler is not None and not self.scheduler.dependency_check_passed:
                    s.set_final_status(FinalStatus.FAILED, "Failed to load the job-type dependency graph (cycle).")
                    break
                if self._registration_timed_out() or self._startup_failed():
                    break
                total = s.total_tracked_tasks()
                if total > 0 and s.num_completed_tracked_tasks() == total:
                    log(f"all {total} tracked tasks completed")
                    break
                wait = 1.0
                if deadline is not None:
                    wait = max(0.01, min(wait, deadline - time.monotonic()))
                self.lock.wait(wait)
            if self.single_node:
                s.set_final_status(FinalStatus.SUCCEEDED if self.preprocess_exit_code == 0 else FinalStatus.FAILED,
                                   None if self.preprocess_exit_code == 0 else
                                   f"Single node job failed with exit code {self.preprocess_exit_code}")
            elif s.final_status != FinalStatus.FAILED:
                s.update_session_status()
            ok = s.final_status == FinalStatus.SUCCEEDED
            if not ok:
                log(f"Tony session failed: {s.final_message}")
            return ok

    def reset(self) -> None:
        """Whole-session restart: kill every container, fresh session, bump the session id."""
        log("resetting for a new session attempt")
        self.mark("session_reset", session=self.session_id + 1)
        self.rm.stop_all(new_attempt=True)
        with self.lock:
            self.first_step.clear()      # the resumed session reports its own first step (time to recover)
            self.hb.reset()
            self.session_id += 1
            self.num_am_retries += 1
            self.task_has_missed_hb = False
            self.untracked_task_failed = False
            self.metrics.clear()
            self.adapter = get_am_adapter(self.framework)
            self.session = None
            self.lock.notify_all()

    def stop(self) -> None:
        self.rm.stop_all()
        self.adapter.destroy()
        # untracked survivors were killed by us: give their completion callbacks a moment
        t_end = time.monotonic() + 3.0
        with self.lock:
            while time.monotonic() < t_end and self.session and any(
                    not t.completed and t.container_id for t in self.session.all_tasks()):
                self.lock.wait(0.05)

    def run(self) -> bool:
        if not self.init():
            self._finish(False, self.final_message or "invalid configuration")
            return False
        self.prepare()
        ok = False
        try:
            while True:
                if os.environ.get(K.TEST_AM_CRASH) and self.num_am_retries == 0:
                    log("TEST_AM_CRASH set: simulating an AM failure")
                    ok = False
                    msg = "AM crashed (fault injection)."
                    if self.max_retries <= 0:
                        self._finish(False, msg)
                        return False
                    self.num_am_retries += 1
                    self.session_id += 1
                    continue
                if os.environ.get(K.TEST_AM_THROW_EXCEPTION_CRASH):
                    raise RuntimeError("TEST_AM_THROW_EXCEPTION_CRASH")
                self.events.emit(ev.application_inited(self.app_id, self.conf.total_instances(), self.host,
                                                       f"container_{self.app_id.split('_', 1)[1]}_01_000001"))
                started = self.start()
                ok = self.monitor() if (started or self.session is not None) else False
                if ok or self.num_am_retries >= self.max_retries or self.single_node or self.client_signal_to_stop:
                    break
                self.reset()
        except Exception as exc:  # noqa: BLE001
            log("AM failed: " + "".join(traceback.format_exception_only(type(exc), exc)).strip())
            traceback.print_exc()
            if self.session is not None:
                self.session.set_final_status(FinalStatus.FAILED, f"AM exception: {exc}")
            ok = False
        self.stop()
        msg = self.session.final_message if self.session else self.final_message
        self._finish(ok, msg)
        return ok

    def _finish(self, ok: bool, message: Optional[str]) -> None:
        with self.lock:
            self.final_status = FinalStatus.SUCCEEDED if ok else FinalStatus.FAILED
            self.final_message = message
            self.app_state = "FINISHED"
            s = self.session
            finished = s.num_completed_tasks() if s else 0
            failed = s.num_failed_tasks() if s else 0
            self.lock.notify_all()
        if self.events is not None:
            self.events.emit(ev.application_finished(self.app_id, finished, failed, {}))
            self.events.stop(self.final_status)
        with open(os.path.join(self.app_dir, "final_status.json.tmp"), "w") as f:
            json.dump({"status": self.final_status, "message": message, "sessionId": self.session_id,
                       "firstStep": self.first_step, "submitTime": self.submit_time,
                       "timeline": self.timeline}, f)
        os.replace(os.path.join(self.app_dir, "final_status.json.tmp"),
                   os.path.join(self.app_dir, "final_status.json"))
        # wait for the client's finishApplication so it can read final task states
        wait_s = self.conf.get_int(K.AM_WAIT_CLIENT_STOP_TIMEOUT_SEC, 15)
        t_end = time.monotonic() + wait_s
        with self.lock:
            while not self.client_signal_to_stop and time.monotonic() < t_end and hasattr(self, "app_rpc"):
                self.lock.wait(0.1)
        if self.hb:
            self.hb.stop()
        if self.rm:
            self.rm.shutdown()
        for srv in ("app_rpc", "metrics_rpc"):
            if hasattr(self, srv):
                getattr(self, srv).stop()
        try:
            with open(os.path.join(self.app_dir, "timeline.json"), "w") as f:
                json.dump(self.timeline, f, indent=1)
        except OSError:
            pass
        log(f"application finished: {self.final_status} ({message})")


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--app_dir", required=True)
    ap.add_argument("--app_id", required=True)
    ap.add_argument("--user", default=None)
    args = ap.parse_args(argv)
    am = ApplicationMaster(args.app_dir, args.app_id, args.user)
    return 0 if am.run() else 1


if __name__ == "__main__":
    sys.exit(main())
