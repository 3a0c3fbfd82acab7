"""Build/version information injected into every job's configuration as ``tony.version-info.*``
(reference: tony-core/src/main/java/com/linkedin/tony/util/VersionInfo.java:133-141, generated
by gradle/version-info.gradle:9-83; here derived from git at run time)."""
from __future__ import annotations

import functools
import os
import subprocess
import time
from typing import Dict

VERSION = "0.1.0"
_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _git(*args: str) -> str:
    try:
        return subprocess.run(["git", "-C", _ROOT, *args], capture_output=True, text=True,
                              timeout=5).stdout.strip()
    except (OSError, subprocess.SubprocessError):
        return ""


@functools.lru_cache(maxsize=1)
def version_info() -> Dict[str, str]:
    return {
        "version": VERSION,
        "revision": _git("rev-parse", "HEAD") or "unknown",
        "branch": _git("rev-parse", "--abbrev-ref", "HEAD") or "unknown",
        "user": os.environ.get("USER", "unknown"),
        "date": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()),
        "url": _git("config", "--get", "remote.origin.url") or "unknown",
        "checksum": (_git("rev-parse", "HEAD:") or "unknown"),
    }
