"""Docker container runtime support.

The reference only sets the YARN docker-runtime environment variables and lets the
NodeManager do the rest (tony-core/src/main/java/com/linkedin/tony/HadoopCompatibleAdapter.java
:95-142; expected values in TestHadoopCompatibleAdapter.java:73-89).  The same variables are
exported here for compatibility, and — since there is no NodeManager — the task command is
additionally wrapped in ``docker run`` (GPUs passed through with ``--gpus device=...``) when a
docker binary exists on the box.
"""
from __future__ import annotations

import os
import shutil
from typing import Dict, List, Optional

from ..conf import keys as K
from ..conf.configuration import TonyConfiguration

ENV_CONTAINER_TYPE = "YARN_CONTAINER_RUNTIME_TYPE"
ENV_DOCKER_IMAGE = "YARN_CONTAINER_RUNTIME_DOCKER_IMAGE"
ENV_DOCKER_MOUNTS = "YARN_CONTAINER_RUNTIME_DOCKER_MOUNTS"


def docker_image(conf: TonyConfiguration, job: str) -> Optional[str]:
    return conf.get(K.docker_image_key(job)) or conf.get(K.DOCKER_CONTAINERS_IMAGE)


def container_runtime_env(conf: TonyConfiguration, job: str) -> Dict[str, str]:
    if not conf.get_bool(K.DOCKER_ENABLED, False):
        return {}
    image = docker_image(conf, job)
    if not image:
        return {}
    env = {ENV_CONTAINER_TYPE: "docker", ENV_DOCKER_IMAGE: image}
    mounts = conf.get(K.DOCKER_CONTAINERS_MOUNT)
    if mounts:
        env[ENV_DOCKER_MOUNTS] = mounts
    return env


def wrap_command(conf: TonyConfiguration, job: str, cmd: List[str], container, container_dir: str) -> List[str]:
    if not conf.get_bool(K.DOCKER_ENABLED, False):
        return cmd
    image = docker_image(conf, job)
    docker = shutil.which("docker")
    if not image or not docker:
        return cmd  # env contract is still exported; nothing to run it with on this box
    out = [docker, "run", "--rm", "--network=host", "--ipc=host", "-v", f"{container_dir}:{container_dir}",
           "-w", container_dir]
    if getattr(container, "gpu_ids", None):
        out += ["--gpus", "device=" + ",".join(str(g) for g in container.gpu_ids)]
    for m in (conf.get(K.DOCKER_CONTAINERS_MOUNT) or "").split(","):
        m = m.strip()
        if m:
            parts = m.split(":")
            out += ["-v", m if len(parts) >= 2 else f"{m}:{m}"]
    out += ["--env-file", os.path.join(container_dir, "docker.env"), image] + cmd
    return out


def write_env_file(conf: TonyConfiguration, env: Dict[str, str], container_dir: str) -> Optional[str]:
    """``docker run --env-file`` input: the task's environment contract, one KEY=VALUE per line."""
    if not conf.get_bool(K.DOCKER_ENABLED, False):
        return None
    os.makedirs(container_dir, exist_ok=True)
    path = os.path.join(container_dir, "docker.env")
    with open(path, "w") as f:
        for k, v in env.items():
            if "\n" not in v:
                f.write(f"{k}={v}\n")
    return path
