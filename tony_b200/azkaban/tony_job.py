"""Workflow-scheduler job type: turn a flat job-properties map into a TonY submission.

Reference: tony-azkaban/src/main/java/com/linkedin/tony/azkaban/TonyJob.java — every
``tony.*`` job property becomes a ``tony.xml`` entry :54-68, flow identity is attached as
application tags, ``worker_env.*`` properties become ``-shell_env`` pairs and the
``executes/task_params/src_dir/python_*`` properties become the matching TonyClient flags
:136-193 (names in TonyJobArg.java:8-15).  Azkaban itself is not required: anything that can
hand over a ``.job``/properties dict (Azkaban, Airflow, cron) can call :func:`run_job`.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Tuple

from ..cli.cluster_submitter import ClusterSubmitter
from ..conf.configuration import write_xml

TONY_CONF_PREFIX = "tony."
WORKER_ENV_PREFIX = "worker_env."
TAG_KEYS = ("azkaban.flow.execid", "azkaban.flow.flowid", "azkaban.flow.projectname", "azkaban.webserver.url")
ARG_PROPS = ("hdfs_classpath", "task_params", "python_binary_path", "python_venv", "executes")
INPUT_DATASET_PROP, OUTPUT_DATASET_PROP = "azkaban.input.dataset", "azkaban.output.dataset"


def read_properties(path: str) -> Dict[str, str]:
    props: Dict[str, str] = {}
    with open(path) as f:
        for line in f:
            line = line.strip()
            if not line or line.startswith(("#", "!")):
                continue
            k, _, v = line.partition("=")
            props[k.strip()] = v.strip()
    return props


def job_configuration(props: Dict[str, str]) -> List[Tuple[str, str, dict]]:
    items = [(k, v, {}) for k, v in props.items() if k.startswith(TONY_CONF_PREFIX)]
    tags = ",".join(f"{k.rsplit('.', 1)[-1]}:{props[k]}" for k in TAG_KEYS if k in props)
    items.append(("tony.application.tags", tags, {}))
    return items


def main_arguments(props: Dict[str, str], conf_file: Optional[str] = None) -> List[str]:
    args: List[str] = []
    if conf_file:
        args += ["-conf_file", conf_file]
    args += ["-src_dir", props.get("src_dir", "src")]
    for k, v in props.items():
        if k.startswith(WORKER_ENV_PREFIX):
            args += ["-shell_env", f"{k[len(WORKER_ENV_PREFIX):]}={v}"]
    for name in ARG_PROPS:
        if props.get(name) is not None:
            args += [f"-{name}", props[name]]
    for prop, env in ((INPUT_DATASET_PROP, "AZKABAN_INPUT_DATASET"), (OUTPUT_DATASET_PROP, "AZKABAN_OUTPUT_DATASET")):
        if props.get(prop) is not None:
            args += ["-shell_env", f"{env}={props[prop].replace(',', ';')}"]
    return args


def run_job(props: Dict[str, str], working_dir: str = ".") -> int:
    os.makedirs(working_dir, exist_ok=True)
    conf_file = os.path.join(working_dir, "_tony-conf-" + props.get("azkaban.job.id", "job"), "tony.xml")
    os.makedirs(os.path.dirname(conf_file), exist_ok=True)
    write_xml(conf_file, job_configuration(props))
    argv = main_arguments(props, conf_file)
    if not os.path.isdir(argv[argv.index("-src_dir") + 1]):
        i = argv.index("-src_dir")
        del argv[i:i + 2]
    return ClusterSubmitter().submit(argv)


def main(argv=None) -> int:
    import sys

    a = sys.argv[1:] if argv is None else argv
    if len(a) != 1:
        sys.stderr.write("usage: python -m tony_b200.azkaban.tony_job <file.job>\n")
        return 2
    return run_job(read_properties(a[0]), os.path.dirname(os.path.abspath(a[0])))


if __name__ == "__main__":
    raise SystemExit(main() & 0xFF)
