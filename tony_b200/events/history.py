"""Job-history files: naming, writing, parsing, and the mover / purger the portal runs.

Layout and rules kept from the reference:
``<history>/intermediate/<appId>/{tony-final.xml, <appId>-<start>-<user>.jhist.inprogress}``
-> ``<appId>-<start>-<end>-<user>-<STATUS>.jhist`` on finish
(util/HistoryFileUtils.java:12-32, events/EventHandler.java:43-60,97-155) -> moved to
``finished/yyyy/MM/dd/<appId>/`` (tony-portal/app/history/HistoryFileMover.java:74-168) ->
purged after the retention period (HistoryFilePurger.java:49-112).  Name validation and
"latest jhist wins" follow util/ParserUtils.java:67-131; metadata/config/event parsing
:153-285; models/JobMetadata.java, JobEvent, JobConfig, JobLog.
"""
from __future__ import annotations

import datetime as _dt
import os
import queue
import re
import shutil
import threading
import time
from dataclasses import dataclass
from typing import List, Optional

from ..conf import keys as K
from ..conf.configuration import read_xml
from . import avro, schema

JHIST = "jhist"
INPROGRESS = "inprogress"
APP_ID_RE = re.compile(r"^application_\d+_\d+$")


# ---- file names ------------------------------------------------------------------------------
def inprogress_name(app_id: str, started_ms: int, user: str) -> str:
    return f"{app_id}-{started_ms}-{user}.{JHIST}.{INPROGRESS}"


def finished_name(app_id: str, started_ms: int, completed_ms: int, user: str, status: str) -> str:
    return f"{app_id}-{started_ms}-{completed_ms}-{user}-{status}.{JHIST}"


def is_valid_hist_file_name(name: str, job_id_regex: str = APP_ID_RE.pattern) -> bool:
    if "." not in name:
        return False
    stem = name[: name.index(".")]
    parts = stem.split("-")
    if len(parts) < 3:
        return False
    if not re.match(job_id_regex, parts[0]) or not parts[1].isdigit():
        return False
    if name.endswith("." + INPROGRESS):
        return len(parts) == 3 and parts[2] == parts[2].lower()
    if not name.endswith("." + JHIST) or len(parts) != 5:
        return False
    return parts[2].isdigit() and parts[3] == parts[3].lower() and parts[4] == parts[4].upper()


def latest_jhist(job_dir: str) -> Optional[str]:
    """Several attempts may leave several files; the greatest start time wins."""
    best, best_start = None, -1
    try:
        names = os.listdir(job_dir)
    except OSError:
        return None
    for n in names:
        if not (n.endswith("." + JHIST) or n.endswith(f".{JHIST}.{INPROGRESS}")):
            continue
        if not is_valid_hist_file_name(n):
            continue
        start = int(n.split("-")[1])
        if start > best_start:
            best, best_start = n, start
    return best


# ---- models ---------------------------------------------------------------------------------
@dataclass
class JobMetadata:
    id: str
    started: int
    completed: int
    status: str
    user: str
    job_link: str = ""
    config_link: str = ""

    @staticmethod
    def from_file_name(name: str) -> Optional["JobMetadata"]:
        if not is_valid_hist_file_name(name):
            return None
        parts = name[: name.index(".")].split("-")
        if len(parts) == 3:
            md = JobMetadata(parts[0], int(parts[1]), -1, "RUNNING", parts[2])
        else:
            md = JobMetadata(parts[0], int(parts[1]), int(parts[2]), parts[4], parts[3])
        md.job_link = f"/jobs/{md.id}"
        md.config_link = f"/config/{md.id}"
        return md


@dataclass
class JobEvent:
    type: str
    event: dict
    timestamp: int


@dataclass
class JobConfig:
    name: str
    value: str
    final: bool = False
    source: str = ""


@dataclass
class JobLog:
    task_type: str
    task_index: int
    host: str
    container_id: str
    url: str


def parse_metadata(job_dir: str) -> Optional[JobMetadata]:
    name = latest_jhist(job_dir)
    return JobMetadata.from_file_name(name) if name else None


def parse_config(job_dir: str) -> List[JobConfig]:
    for cand in (K.TONY_FINAL_XML, "config.xml"):
        p = os.path.join(job_dir, cand)
        if os.path.exists(p):
            return [JobConfig(n, v, a.get("final", "false") == "true", a.get("source", ""))
                    for n, v, a in read_xml(p)]
    return []


def parse_events(job_dir: str) -> List[JobEvent]:
    name = latest_jhist(job_dir)
    if not name:
        return []
    try:
        _, objs = avro.read_container(os.path.join(job_dir, name))
    except (OSError, ValueError, EOFError, KeyError):
        return []
    return [JobEvent(o["type"], o["event"], o["timestamp"]) for o in objs]


def parse_logs(job_dir: str, user: str = "") -> List[JobLog]:
    """One log link per started task (reference builds NodeManager URLs, models/JobLog.java:29,
    154-168; here the container's stdout/stderr live on the local file system)."""
    logs = []
    for e in parse_events(job_dir):
        if e.type != "TASK_STARTED":
            continue
        ev = e.event
        cid = ev.get("containerID") or ""
        logs.append(JobLog(ev["taskType"], ev["taskIndex"], ev["host"], cid,
                           f"file://{job_dir}/containers/{cid}/stdout" if cid else ""))
    return logs


# ---- writer ---------------------------------------------------------------------------------
class EventHandler(threading.Thread):
    """Drains a queue of events into the in-progress jhist; ``stop`` renames it to its final name."""

    def __init__(self, job_dir: str, app_id: str, started_ms: int, user: str):
        super().__init__(name="event-handler", daemon=True)
        self.job_dir, self.app_id, self.started_ms, self.user = job_dir, app_id, started_ms, user
        os.makedirs(job_dir, exist_ok=True)
        self.path = os.path.join(job_dir, inprogress_name(app_id, started_ms, user))
        self._q: "queue.Queue[Optional[dict]]" = queue.Queue()
        self._writer = avro.DataFileWriter(self.path, schema.EVENT)
        self.final_path: Optional[str] = None

    def emit(self, event: dict) -> None:
        self._q.put(event)

    def run(self) -> None:
        while True:
            ev = self._q.get()
            if ev is None:
                return
            try:
                self._writer.append(ev)
                self._writer.flush()
            except Exception:  # noqa: BLE001 - history must never take the job down
                pass

    def stop(self, status: str, completed_ms: Optional[int] = None) -> str:
        self._q.put(None)
        if self.is_alive():
            self.join(timeout=10)
        self._writer.close()
        completed_ms = completed_ms or int(time.time() * 1000)
        self.final_path = os.path.join(
            self.job_dir, finished_name(self.app_id, self.started_ms, completed_ms, self.user, status))
        os.replace(self.path, self.final_path)
        return self.final_path


# ---- mover / purger ---------------------------------------------------------------------------
def finished_subdir(completed_ms: int, tz_name: str = "UTC") -> str:
    try:
        from zoneinfo import ZoneInfo

        tz = ZoneInfo(tz_name)
    except Exception:  # noqa: BLE001
        tz = _dt.timezone.utc
    d = _dt.datetime.fromtimestamp(completed_ms / 1000.0, tz)
    return f"{d.year:04d}/{d.month:02d}/{d.day:02d}"


class HistoryFileMover:
    def __init__(self, intermediate: str, finished: str, tz_name: str = "UTC",
                 killed_apps=lambda: set(), on_update=lambda job_dir: None):
        self.intermediate, self.finished, self.tz_name = intermediate, finished, tz_name
        self.killed_apps = killed_apps
        self.on_update = on_update

    def rename_killed(self, job_dir: str, app_id: str) -> None:
        name = latest_jhist(job_dir)
        if not name or not name.endswith("." + INPROGRESS):
            return
        parts = name[: name.index(".")].split("-")
        new = finished_name(parts[0], int(parts[1]), int(time.time() * 1000), parts[2], "KILLED")
        os.replace(os.path.join(job_dir, name), os.path.join(job_dir, new))

    def run_once(self) -> List[str]:
        moved = []
        if not os.path.isdir(self.intermediate):
            return moved
        killed = set(self.killed_apps())
        for app_id in sorted(os.listdir(self.intermediate)):
            job_dir = os.path.join(self.intermediate, app_id)
            if not os.path.isdir(job_dir):
                continue
            if app_id in killed:
                self.rename_killed(job_dir, app_id)
            self.on_update(job_dir)
            name = latest_jhist(job_dir)
            if not name or name.endswith("." + INPROGRESS):
                continue
            md = JobMetadata.from_file_name(name)
            dest_parent = os.path.join(self.finished, finished_subdir(md.completed, self.tz_name))
            os.makedirs(dest_parent, exist_ok=True)
            dest = os.path.join(dest_parent, app_id)
            if os.path.exists(dest):
                shutil.rmtree(dest)
            shutil.move(job_dir, dest)
            moved.append(dest)
        return moved


class HistoryFilePurger:
    def __init__(self, finished: str, retention_sec: int):
        self.finished, self.retention_sec = finished, retention_sec

    def run_once(self, now: Optional[float] = None) -> List[str]:
        """Delete ``yyyy/MM/dd`` directories older than the retention period."""
        now = time.time() if now is None else now
        cutoff = _dt.datetime.fromtimestamp(now - self.retention_sec, _dt.timezone.utc).date()
        removed = []
        if not os.path.isdir(self.finished):
            return removed
        for y in sorted(os.listdir(self.finished)):
            for m in sorted(os.listdir(os.path.join(self.finished, y))) if y.isdigit() else []:
                mdir = os.path.join(self.finished, y, m)
                for d in sorted(os.listdir(mdir)) if m.isdigit() else []:
                    if not d.isdigit():
                        continue
                    if _dt.date(int(y), int(m), int(d)) < cutoff:
                        shutil.rmtree(os.path.join(mdir, d), ignore_errors=True)
                        removed.append(os.path.join(mdir, d))
                if m.isdigit() and not os.listdir(mdir):
                    os.rmdir(mdir)
            ydir = os.path.join(self.finished, y)
            if y.isdigit() and not os.listdir(ydir):
                os.rmdir(ydir)
        return removed


def find_job_dir(history_root_finished: str, intermediate: str, app_id: str) -> Optional[str]:
    cand = os.path.join(intermediate, app_id)
    if os.path.isdir(cand):
        return cand
    for root, dirs, _ in os.walk(history_root_finished):
        if app_id in dirs:
            return os.path.join(root, app_id)
    return None
