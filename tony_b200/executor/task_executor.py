"""TaskExecutor: the per-container agent that registers with the AM and runs the user command.

Lifecycle parity with tony-core/src/main/java/com/linkedin/tony/TaskExecutor.java:
initConfigs :240-281 -> extract src/venv :158 -> metrics thread :166-170 -> reserve rpc (+TB)
port :89-101 -> heartbeater + registerWorkerSpec until the gang barrier opens :283-297 ->
release ports unless *_REUSE_PORT :205-215 -> framework adapter run :219 ->
registerExecutionResult :223,313-320 -> exit with the user's code :237.  Fault-injection hooks
TEST_TASK_EXECUTOR_NUM_HB_MISS :326-349 and TEST_TASK_EXECUTOR_SKEW :364-384 keep their names.

Differences by design: ``registerWorkerSpec`` is a long-poll (the AM answers the instant the
barrier opens; the reference re-polls every 3 s), and a watcher reports the user runtime's
first-step marker so submit-to-first-step latency is measured.
"""
from __future__ import annotations

import os
import subprocess
import sys
import threading
import time
from typing import Dict, List, Optional

from ..conf import keys as K
from ..conf.configuration import TonyConfiguration
from ..rpc.server import RpcClient
from ..runtimes.frameworks import get_task_adapter
from ..utils import utils as U
from .ports import ServerPort, reserve
from .task_monitor import TaskMonitor

MAX_CONSECUTIVE_HB_FAILURES = 5


class Heartbeater(threading.Thread):
    def __init__(self, client: RpcClient, task_id: str, interval_ms: int, num_hb_miss: int = 0):
        super().__init__(name="heartbeater", daemon=True)
        self.client, self.task_id = client, task_id
        self.interval_s = interval_ms / 1000.0
        self.num_hb_miss = num_hb_miss  # test hook: skip this many beats once
        self._stop = threading.Event()
        self.failures = 0

    def run(self) -> None:
        while not self._stop.wait(self.interval_s):
            if self.num_hb_miss > 0:
                self.num_hb_miss -= 1
                continue
            try:
                self.client.call("taskExecutorHeartbeat", taskId=self.task_id)
                self.failures = 0
            except Exception:  # noqa: BLE001
                self.failures += 1
                if self.failures >= MAX_CONSECUTIVE_HB_FAILURES:
                    sys.stderr.write("Exceeded max consecutive heartbeat failures; giving up.\n")
                    os._exit(1)

    def stop(self) -> None:
        self._stop.set()


class TaskExecutor:
    def __init__(self, env: Optional[Dict[str, str]] = None):
        e = dict(os.environ if env is None else env)
        self.env = e
        self.job_name = e[K.JOB_NAME]
        self.task_index = int(e[K.TASK_INDEX])
        self.num_tasks = int(e.get(K.TASK_NUM, "1"))
        self.is_chief = e.get(K.IS_CHIEF, "false").lower() == "true"
        self.session_id = int(e.get(K.SESSION_ID, "0"))
        self.distributed_mode = e.get(K.DISTRIBUTED_MODE_NAME, "GANG")
        self.am_host = e[K.AM_HOST]
        self.am_port = int(e[K.AM_PORT])
        self.metrics_port = int(e.get(K.METRICS_RPC_PORT, "0"))
        self.conf = TonyConfiguration.from_final_xml(e[K.TONY_CONF_PATH])
        secret = e.get(K.TONY_SECRET)
        self.secret = bytes.fromhex(secret) if secret else None
        self.task_id = U.task_id(self.job_name, self.task_index)
        self.framework = self.conf.get(K.FRAMEWORK, "tensorflow")
        self.shell_env: Dict[str, str] = U.parse_key_value(self.conf.get_strings(K.EXECUTION_ENVS))
        self.task_command = self.conf.get(K.command_key(self.job_name)) or self.conf.get(K.CONTAINERS_COMMAND)
        self.timeout_ms = self.conf.get_int(K.WORKER_TIMEOUT, 0)
        self.hb_interval_ms = self.conf.get_int(K.HEARTBEAT_INTERVAL_MS, 1000)
        self.cluster_spec: Optional[str] = None
        self.rpc_port: Optional[ServerPort] = None
        self.tb_port: Optional[ServerPort] = None
        self.client = RpcClient(self.am_host, self.am_port, self.secret)
        self.hb_client = RpcClient(self.am_host, self.am_port, self.secret, retries=1)
        self.metrics_client = RpcClient(self.am_host, self.metrics_port, self.secret, retries=1) \
            if self.metrics_port else None
        self.monitor: Optional[TaskMonitor] = None
        self.heartbeater: Optional[Heartbeater] = None
        self._children: List[subprocess.Popen] = []
        self.mark_dir = os.path.join(os.getcwd(), ".tony_marks")

    # -- helpers used by the framework adapters ---------------------------------------------
    def register_child(self, proc: subprocess.Popen) -> None:
        self._children.append(proc)

    def callback_info_to_am(self, info: str) -> None:
        self.client.call("registerCallbackInfo", taskId=self.task_id, callbackInfo=info)

    def execute_user_command(self) -> int:
        if not self.task_command:
            sys.stderr.write(f"No command configured for job type {self.job_name}\n")
            return 1
        env = dict(self.shell_env)
        env["TONY_B200_MARK_DIR"] = self.mark_dir
        threading.Thread(target=self._watch_first_step, name="first-step", daemon=True).start()
        return U.execute_shell(self.task_command, self.timeout_ms, env=env, on_start=self.register_child)

    def _watch_first_step(self) -> None:
        path = os.path.join(self.mark_dir, f"first_step.{self.job_name}.{self.task_index}")
        while True:
            if os.path.exists(path):
                try:
                    with open(path) as f:
                        ts = float(f.readline().strip())
                    self.hb_client.call("markFirstStep", taskId=self.task_id, timestamp=ts)
                except Exception:  # noqa: BLE001
                    pass
                return
            time.sleep(0.02)

    # -- lifecycle --------------------------------------------------------------------------
    def extract_resources(self) -> None:
        cwd = os.getcwd()
        for name in os.listdir(cwd):
            if name.startswith(K.TONY_SRC_ZIP_PREFIX) and name.endswith(".zip"):
                U.unzip_archive(os.path.join(cwd, name), cwd)
        venv = os.path.join(cwd, K.PYTHON_VENV_ZIP)
        if os.path.exists(venv) and not os.path.isdir(os.path.join(cwd, K.PYTHON_VENV_DIR)):
            U.unzip_archive(venv, os.path.join(cwd, K.PYTHON_VENV_DIR))

    def start_monitor(self) -> None:
        if self.metrics_client is None:
            return
        gpus = [int(x) for x in os.environ.get("TONY_B200_GPU_IDS", "").split(",") if x.strip().isdigit()]

        def push(metrics: Dict[str, float]) -> None:
            self.metrics_client.call("updateMetrics", taskType=self.job_name, taskIndex=self.task_index,
                                     metrics=[{"name": k, "value": v} for k, v in metrics.items()])

        self.monitor = TaskMonitor(push, self.conf.get_int(K.METRICS_INTERVAL_MS, 5000),
                                   gpu_enabled=self.conf.get_bool(K.GPU_METRICS_ENABLED, True) and bool(gpus),
                                   gpu_exec_path=self.conf.get(K.GPU_EXEC_PATH), visible_gpus=gpus or None,
                                   user_metrics_path=os.path.join(self.mark_dir,
                                                                  f"metrics.{self.job_name}.{self.task_index}"))
        self.monitor.start()

    def setup_ports(self, adapter) -> None:
        self.rpc_port = reserve(self.env.get("TF_GRPC_REUSE_PORT", "false").lower() == "true")
        if adapter.need_reserve_tb_port():
            self.tb_port = reserve(self.env.get("TB_SERVER_REUSE_PORT", "false").lower() == "true")
            self.shell_env[K.TB_PORT] = str(self.tb_port.port)
            self.client.call("registerTensorBoardUrl", spec=f"{U.local_hostname()}:{self.tb_port.port}")

    def register_and_get_cluster_spec(self) -> str:
        miss = int(self.env.get(K.TEST_TASK_EXECUTOR_NUM_HB_MISS, "0") or 0)
        self.heartbeater = Heartbeater(self.hb_client, self.task_id, self.hb_interval_ms, miss)
        self.heartbeater.start()
        host_port = f"{U.local_hostname()}:{self.rpc_port.port}"
        while True:
            spec = self.client.call("registerWorkerSpec", worker=self.task_id, spec=host_port)
            if spec is not None:
                return spec

    def release_ports(self) -> None:
        for p, flag in ((self.rpc_port, "TF_GRPC_REUSE_PORT"), (self.tb_port, "TB_SERVER_REUSE_PORT")):
            if p is not None and self.env.get(flag, "false").lower() != "true":
                p.close()

    def skew_if_requested(self) -> None:
        """TEST_TASK_EXECUTOR_SKEW=jobtype#index#ms makes that one task a straggler."""
        spec = self.env.get(K.TEST_TASK_EXECUTOR_SKEW)
        if not spec:
            return
        try:
            job, idx, ms = spec.split("#")
            if job == self.job_name and int(idx) == self.task_index:
                time.sleep(int(ms) / 1000.0)
        except ValueError:
            pass

    def run(self) -> int:
        self.extract_resources()
        self.start_monitor()
        adapter = get_task_adapter(self.framework, self)
        self.setup_ports(adapter)
        self.cluster_spec = self.register_and_get_cluster_spec()
        self.release_ports()
        try:
            exit_code = adapter.run()
        except Exception as exc:  # noqa: BLE001
            sys.stderr.write(f"task {self.task_id} failed before/while running the user command: {exc}\n")
            exit_code = 1
        finally:
            for p in (self.rpc_port, self.tb_port):
                if p is not None:
                    p.close()
        self.skew_if_requested()
        if self.monitor:
            self.monitor.flush()      # before the result: the AM emits TaskFinished on container exit
        try:
            self.client.call("registerExecutionResult", exitCode=exit_code, jobName=self.job_name,
                             jobIndex=str(self.task_index), sessionId=str(self.session_id))
        except Exception as exc:  # noqa: BLE001
            sys.stderr.write(f"could not register execution result: {exc}\n")
        if self.monitor:
            self.monitor.stop()
        if self.heartbeater:
            self.heartbeater.stop()
        for c in self._children:
            if c.poll() is None:
                U.kill_process_group(c)
        return exit_code


def main() -> int:
    return TaskExecutor().run()


if __name__ == "__main__":
    sys.exit(main())
