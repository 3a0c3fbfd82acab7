"""TonyClient: parse the CLI, freeze the configuration, stage resources, launch and watch the AM.

Public surface kept from tony-core/src/main/java/com/linkedin/tony/TonyClient.java:
``init(argv)`` :442, ``start()`` :1290 (returns 0 / -1), ``force_kill_application()`` :1268,
``add_listener(TaskUpdateListener)`` :1388, ``CallbackHandler.on_application_id_received``
(client/CallbackHandler.java:13-16), ``get_tony_conf()`` :859; CLI flags :425-436 with
GnuParser's ``-x``/``--x``/``--x=v`` forms; conf layering :657-691; limit validation :788-857;
task-command synthesis :618-650; sidecar TensorBoard :559-600; monitor loop :1031-1111.

Re-designed for the local cluster: the "submission" is a detached ApplicationMaster process,
resources are staged into ``<work-dir>/<appId>/staging`` instead of HDFS, and the monitor
polls the AM every 100 ms (the reference sleeps 1 s per round).
"""
from __future__ import annotations

import fcntl
import json
import os
import secrets
import shutil
import subprocess
import sys
import time
from typing import Dict, List, Optional, Set

from ..am.session import TaskInfo
from ..conf import keys as K
from ..conf.configuration import TonyConfiguration, build_configuration, parse_memory_string
from ..rpc.server import RpcClient
from ..utils import utils as U
from ..utils.localizable_resource import stage_resource
from ..version import version_info

REPO_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
SIDECAR_TB_SCRIPT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                 "executor", "sidecar_tensorboard.py")

# name -> takes a value, repeatable
OPTIONS = {
    "executes": (True, False), "task_params": (True, False), "shell_env": (True, True),
    "container_env": (True, True), "conf": (True, True), "conf_file": (True, False),
    "src_dir": (True, False), "sidecar_tensorboard_log_dir": (True, False), "help": (False, False),
    "hdfs_classpath": (True, False), "python_binary_path": (True, False), "python_venv": (True, False),
}


class UsageError(ValueError):
    pass


def parse_cli(argv: List[str]) -> Dict[str, object]:
    """GnuParser semantics: ``-opt v``, ``--opt v`` and ``--opt=v`` are all accepted."""
    out: Dict[str, object] = {}
    i = 0
    while i < len(argv):
        tok = argv[i]
        if not tok.startswith("-"):
            raise UsageError(f"unexpected argument {tok!r}")
        name = tok.lstrip("-")
        value: Optional[str] = None
        if "=" in name and name.split("=", 1)[0] in OPTIONS:
            name, value = name.split("=", 1)
        if name not in OPTIONS:
            raise UsageError(f"Unrecognized option: {tok}")
        takes_value, repeatable = OPTIONS[name]
        if takes_value:
            if value is None:
                i += 1
                if i >= len(argv):
                    raise UsageError(f"Missing argument for option: {name}")
                value = argv[i]
            if repeatable:
                out.setdefault(name, []).append(value)  # type: ignore[union-attr]
            else:
                out[name] = value
        else:
            out[name] = True
        i += 1
    return out


def usage() -> str:
    lines = ["usage: TonyClient"]
    for name, (takes, rep) in OPTIONS.items():
        lines.append(f"  -{name}{' <arg>' if takes else ''}{' (repeatable)' if rep else ''}")
    return "\n".join(lines)


class TaskUpdateListener:
    def on_task_infos_updated(self, task_infos: Set[TaskInfo]) -> None:  # pragma: no cover - interface
        raise NotImplementedError


class CallbackHandler:
    def on_application_id_received(self, app_id: str) -> None:  # pragma: no cover - interface
        raise NotImplementedError


def validate_tony_conf(conf: TonyConfiguration) -> None:
    """Instance and resource limits (TonyClient.java:788-857); raises RuntimeError."""
    requests = U.parse_container_requests(conf)
    for job, req in requests.items():
        limit = conf.get_int(K.max_instances_key(job), -1)
        if 0 <= limit < req.num_instances:
            raise RuntimeError(f"Job requested {req.num_instances} {job} task instances but the limit is "
                               f"{limit} {job} task instances.")
    max_total = conf.get_int(K.MAX_TOTAL_INSTANCES, -1)
    total = sum(r.num_instances for r in requests.values())
    if 0 <= max_total < total:
        raise RuntimeError(f"Job requested {total} total task instances but limit is {max_total}.")
    for key, value in list(conf.items()):
        m = K.MAX_TOTAL_RESOURCES_REGEX.match(key)
        if not m or m.group(1) == "instances":
            continue
        limit = int(value)
        if limit < 0:
            continue
        resource = m.group(1)
        requested = 0
        for job in conf.job_types():
            v = conf.get(K.resource_key(job, resource))
            if v is not None:
                per_task = parse_memory_string(v) if resource == "memory" else int(v)
                requested += per_task * conf.num_instances(job)
        if requested > limit:
            raise RuntimeError(f"Total amount of {resource} ({requested}) requested exceeds maximum allowed of {limit}")


class TonyClient:
    def __init__(self, conf: Optional[TonyConfiguration] = None, callback_handler: Optional[CallbackHandler] = None):
        self.base_conf = conf
        self.callback_handler = callback_handler
        self.listeners: List[TaskUpdateListener] = []
        self.tony_conf: Optional[TonyConfiguration] = None
        self.app_id: Optional[str] = None
        self.app_dir: Optional[str] = None
        self.am_proc: Optional[subprocess.Popen] = None
        self.rpc: Optional[RpcClient] = None
        self.task_infos: Set[TaskInfo] = set()
        self.final_status: Optional[str] = None
        self.final_message: Optional[str] = None
        self.submit_to_first_step_ms: Optional[float] = None
        self.src_dir: Optional[str] = None
        self.python_venv: Optional[str] = None
        self.quiet = os.environ.get("TONY_B200_QUIET") == "1"

    # ------------------------------------------------------------------ configuration
    def init(self, argv: List[str]) -> bool:
        cli = parse_cli(list(argv))
        if cli.get("help"):
            print(usage())
            return False
        conf = build_configuration(cli.get("conf_file"), cli.get("conf", []))  # type: ignore[arg-type]
        if self.base_conf is not None:
            merged = self.base_conf.copy()
            for k, v in conf.items():
                if conf.source_of(k) != K.TONY_DEFAULT_XML or k not in merged:
                    merged.set(k, v, conf.source_of(k) or "")
            conf = merged
        for k, v in version_info().items():
            conf.set(K.VERSION_INFO_PREFIX + k, v, "version-info")
        self.src_dir = cli.get("src_dir")  # type: ignore[assignment]
        self.python_venv = cli.get("python_venv")  # type: ignore[assignment]
        python_binary = cli.get("python_binary_path")
        cmd = U.build_task_command(self.python_venv, python_binary, cli.get("executes"), cli.get("task_params"))
        if cmd is not None:
            conf.set(K.CONTAINERS_COMMAND, cmd, "cli")
        if python_binary:
            conf.set(K.PYTHON_EXEC_PATH, python_binary, "cli")
        shell_env = list(cli.get("shell_env", []))  # type: ignore[arg-type]
        for kv in shell_env:
            conf.set_cli(K.EXECUTION_ENVS, kv)
        for kv in cli.get("container_env", []):  # type: ignore[union-attr]
            conf.set_cli(K.CONTAINERS_ENVS, kv)
        tb_dir = cli.get("sidecar_tensorboard_log_dir") or conf.get(K.TENSORBOARD_LOG_DIR)
        if tb_dir:
            self._set_sidecar_tb(conf, str(tb_dir), python_binary)
        validate_tony_conf(conf)
        if conf.total_instances() == 0 and not (conf.get(K.AM_COMMAND) or conf.get(K.CONTAINERS_COMMAND)):
            raise UsageError("nothing to run: no task instances and no command (-executes / tony.am.command)")
        self.tony_conf = conf
        return True

    def _set_sidecar_tb(self, conf: TonyConfiguration, log_dir: str, python_binary) -> None:
        tb = K.TENSORBOARD_JOB_NAME
        conf.set(K.TENSORBOARD_LOG_DIR, log_dir, "cli")
        conf.set(K.instances_key(tb), "1", "sidecar-tb")
        conf.set(K.vcores_key(tb), conf.get(K.vcores_key(tb), "2"), "sidecar-tb")
        conf.set(K.memory_key(tb), conf.get(K.memory_key(tb), "2g"), "sidecar-tb")
        conf.set(K.gpus_key(tb), conf.get(K.gpus_key(tb), "0"), "sidecar-tb")
        sidecars = conf.sidecar_job_types()
        if tb not in sidecars:
            sidecars.append(tb)
        conf.set(K.SIDECAR_JOBTYPES, ",".join(sidecars), "sidecar-tb")
        if not conf.get(K.command_key(tb)):
            interp = python_binary or sys.executable
            conf.set(K.command_key(tb), f"{interp} {os.path.basename(SIDECAR_TB_SCRIPT)}", "sidecar-tb")
        conf.set_cli(K.resources_key(tb), SIDECAR_TB_SCRIPT)
        conf.set_cli(K.EXECUTION_ENVS, f"{K.TB_LOG_DIR}={log_dir}")
        if os.environ.get("SIDECAR_TB_TEST"):
            conf.set_cli(K.EXECUTION_ENVS, "SIDECAR_TB_TEST=true")

    def get_tony_conf(self) -> TonyConfiguration:
        return self.tony_conf

    def add_listener(self, listener: TaskUpdateListener) -> None:
        self.listeners.append(listener)

    # ------------------------------------------------------------------ submission
    def work_dir(self) -> str:
        d = (self.tony_conf.get(K.B200_WORK_DIR) if self.tony_conf else None) or \
            os.environ.get("TONY_B200_WORK_DIR") or os.path.join(os.path.expanduser("~"), ".tony_b200")
        os.makedirs(d, exist_ok=True)
        return d

    def _new_application_id(self) -> str:
        cluster = os.path.join(self.work_dir(), "cluster")
        os.makedirs(cluster, exist_ok=True)
        with open(os.path.join(cluster, "app_seq"), "a+") as f:
            fcntl.flock(f, fcntl.LOCK_EX)
            f.seek(0)
            parts = f.read().split()
            ts = int(parts[0]) if len(parts) == 2 else int(time.time() * 1000)
            seq = (int(parts[1]) if len(parts) == 2 else 0) + 1
            f.seek(0)
            f.truncate()
            f.write(f"{ts} {seq}")
            f.flush()
            fcntl.flock(f, fcntl.LOCK_UN)
        return f"application_{ts}_{seq:04d}"

    def _process_final_conf(self) -> None:
        """Stage src / venv / resources and write tony-final.xml (TonyClient.java:229-311)."""
        conf = self.tony_conf
        staging = os.path.join(self.app_dir, "staging")
        os.makedirs(staging, exist_ok=True)
        extra: List[str] = []
        if self.src_dir:
            z = os.path.join(staging, f"{K.TONY_SRC_ZIP_PREFIX}{self.app_id}.zip")
            U.zip_folder(self.src_dir, z)
            extra.append(z)
        if self.python_venv:
            dst = os.path.join(staging, K.PYTHON_VENV_ZIP)
            shutil.copy2(self.python_venv, dst)
            extra.append(dst)
        keys = [K.CONTAINERS_RESOURCES] + [K.resources_key(j) for j in conf.job_types()]
        for key in keys:
            staged = [stage_resource(r, os.path.join(staging, key.replace(".", "_")))
                      for r in conf.get_strings(key)]
            if key == K.CONTAINERS_RESOURCES:
                staged += extra
            if staged:
                conf.set(key, ",".join(staged), conf.source_of(key) or "client")
        conf.write_xml(os.path.join(self.app_dir, K.TONY_FINAL_XML))

    def _submit(self) -> None:
        conf = self.tony_conf
        env = dict(os.environ)
        env["PYTHONPATH"] = REPO_ROOT + os.pathsep + env.get("PYTHONPATH", "")
        env["TONY_B200_SUBMIT_TIME"] = f"{self.submit_time:.6f}"
        if conf.get_bool(K.SECURITY_ENABLED, True):
            self.secret = secrets.token_bytes(32)
            env[K.TONY_SECRET] = self.secret.hex()
        else:
            self.secret = None
        out = open(os.path.join(self.app_dir, "amstdout.log"), "ab")
        err = open(os.path.join(self.app_dir, "amstderr.log"), "ab")
        am_cmd = [sys.executable, "-m", "tony_b200.am.application_master", "--app_dir", self.app_dir,
                  "--app_id", self.app_id, "--user", U.current_user()]
        if os.environ.get("TEST_AM_NEVER_STARTS"):
            # fault injection: the AM container dies before it publishes an RPC endpoint
            # (the client then has no RPC proxy at all: TestTonyE2E.testNullAMRpcClient)
            am_cmd = [sys.executable, "-c", "import sys; sys.exit(3)"]
        self.am_proc = subprocess.Popen(
            am_cmd,
            env=env, cwd=self.app_dir, stdout=out, stderr=err, start_new_session=True)
        out.close()
        err.close()

    # ------------------------------------------------------------------ monitoring
    def _connect(self) -> bool:
        path = os.path.join(self.app_dir, "am.json")
        if not os.path.exists(path):
            return False
        with open(path) as f:
            info = json.load(f)
        self.rpc = RpcClient(info["host"], info["port"], self.secret, retries=3, retry_interval_s=0.2)
        return True

    def _update_task_infos(self) -> None:
        try:
            rows = self.rpc.call("getTaskInfos")
        except Exception:  # noqa: BLE001
            return
        infos = {TaskInfo.from_dict(r) for r in rows or []}
        changed = {(t.name, t.index, t.status, t.url) for t in infos} != \
            {(t.name, t.index, t.status, t.url) for t in self.task_infos}
        if infos and changed:
            self.task_infos = infos
            for l in self.listeners:
                l.on_task_infos_updated(set(infos))
            if not self.quiet:
                self.print_task_table()

    def print_task_table(self) -> None:
        for t in sorted(self.task_infos, key=lambda t: t.sort_key()):
            print(f"  {t.name}:{t.index:<3} {t.status.name:<10} {t.url}")

    def _monitor(self) -> bool:
        timeout_ms = self.tony_conf.get_int(K.APPLICATION_TIMEOUT, 0)
        deadline = None if timeout_ms <= 0 else time.monotonic() + timeout_ms / 1000.0 + 5.0
        while True:
            if self.rpc is None:
                self._connect()
            status = None
            if self.rpc is not None:
                self._update_task_infos()
                try:
                    status = self.rpc.call("getApplicationStatus")
                except Exception:  # noqa: BLE001
                    status = None
            if status and status.get("submitToFirstStepMs") is not None:
                self.submit_to_first_step_ms = status["submitToFirstStepMs"]
            if status and status.get("state") == "FINISHED":
                self._update_task_infos()
                self.final_status, self.final_message = status["finalStatus"], status.get("message")
                try:
                    self.rpc.call("finishApplication")
                except Exception:  # noqa: BLE001
                    pass
                break
            if self.am_proc.poll() is not None:
                self._read_final_file()
                break
            if deadline is not None and time.monotonic() > deadline:
                if not self.quiet:
                    print("Application timed out; killing it.")
                self.force_kill_application()
                self.final_status, self.final_message = "KILLED", "client-side application timeout"
                break
            time.sleep(0.1)
        try:
            self.am_proc.wait(timeout=30)
        except subprocess.TimeoutExpired:
            self.force_kill_application()
        if self.final_status is None:
            self._read_final_file()
        return self.final_status == "SUCCEEDED"

    def _read_final_file(self) -> None:
        try:
            with open(os.path.join(self.app_dir, "final_status.json")) as f:
                d = json.load(f)
            self.final_status, self.final_message = d["status"], d.get("message")
            fs = d.get("firstStep") or {}
            if fs and self.submit_to_first_step_ms is None:
                self.submit_to_first_step_ms = 1000.0 * (min(fs.values()) - d.get("submitTime", self.submit_time))
        except (OSError, ValueError, KeyError):
            self.final_status = self.final_status or "FAILED"
            self.final_message = self.final_message or "ApplicationMaster exited without a final status"

    def force_kill_application(self) -> None:
        if self.am_proc is not None and self.am_proc.poll() is None:
            # the AM kills its containers on SIGTERM-less exit via process groups; be thorough
            try:
                if self.rpc is not None:
                    self.rpc.call("finishApplication")
            except Exception:  # noqa: BLE001
                pass
            try:
                self.am_proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                U.kill_process_group(self.am_proc)

    # ------------------------------------------------------------------ entry points
    def run(self) -> bool:
        self.submit_time = time.time()
        self.app_id = self._new_application_id()
        self.app_dir = os.path.join(self.work_dir(), self.app_id)
        os.makedirs(self.app_dir, exist_ok=True)
        if self.callback_handler is not None:
            self.callback_handler.on_application_id_received(self.app_id)
        self._process_final_conf()
        self._submit()
        if not self.quiet:
            print(f"Submitted {self.app_id}; logs under {self.app_dir}")
        ok = self._monitor()
        if not self.quiet:
            print(f"Application {self.app_id} finished with status {self.final_status}"
                  + (f": {self.final_message}" if self.final_message else ""))
            if self.submit_to_first_step_ms is not None:
                print(f"submit-to-first-step latency: {self.submit_to_first_step_ms:.0f} ms")
        return ok

    def start(self) -> int:
        try:
            return 0 if self.run() else -1
        except Exception as exc:  # noqa: BLE001
            sys.stderr.write(f"Failed to run TonyClient: {exc}\n")
            return -1


def main(argv=None) -> int:
    client = TonyClient()
    try:
        if not client.init(sys.argv[1:] if argv is None else argv):
            return 0
    except (UsageError, RuntimeError, ValueError) as exc:
        sys.stderr.write(f"{exc}\n{usage()}\n")
        return -1
    return client.start()


if __name__ == "__main__":
    sys.exit(main())
