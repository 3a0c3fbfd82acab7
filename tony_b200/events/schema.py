"""The job-history event schema (byte-compatible with the reference's Avro IDL files
tony-core/src/main/avro/{Event,EventType,ApplicationInited,ApplicationFinished,TaskStarted,
TaskFinished,Metric}.avsc) and plain-dict constructors for each event."""
from __future__ import annotations

import time
from typing import Dict, List, Optional

NS = "com.linkedin.tony.events"

METRIC = {"type": "record", "name": "Metric", "namespace": NS, "fields": [
    {"name": "name", "type": "string"}, {"name": "value", "type": "double"}]}

EVENT_TYPE = {"type": "enum", "name": "EventType", "namespace": NS,
              "symbols": ["APPLICATION_INITED", "APPLICATION_FINISHED", "TASK_STARTED", "TASK_FINISHED"]}

APPLICATION_INITED = {"type": "record", "name": "ApplicationInited", "namespace": NS, "fields": [
    {"name": "applicationId", "type": "string"},
    {"name": "numTasks", "type": "int"},
    {"name": "host", "type": "string"},
    {"name": "containerID", "type": ["null", "string"], "doc": "Container  id", "default": None}]}

APPLICATION_FINISHED = {"type": "record", "name": "ApplicationFinished", "namespace": NS, "fields": [
    {"name": "applicationId", "type": "string"},
    {"name": "finishedTasks", "type": "int"},
    {"name": "failedTasks", "type": "int"},
    {"name": "metrics", "type": {"type": "array", "items": METRIC}}]}

TASK_STARTED = {"type": "record", "name": "TaskStarted", "namespace": NS, "fields": [
    {"name": "taskType", "type": "string"},
    {"name": "taskIndex", "type": "int"},
    {"name": "host", "type": "string"},
    {"name": "containerID", "type": ["null", "string"], "doc": "Container  id", "default": None}]}

TASK_FINISHED = {"type": "record", "name": "TaskFinished", "namespace": NS, "fields": [
    {"name": "taskType", "type": "string"},
    {"name": "taskIndex", "type": "int"},
    {"name": "status", "type": "string"},
    {"name": "metrics", "type": {"type": "array", "items": "Metric"}},
    {"name": "containerDiagnostic", "type": ["null", "string"], "doc": "Container error information",
     "default": None}]}

EVENT = {"type": "record", "name": "Event", "namespace": NS, "fields": [
    {"name": "type", "type": EVENT_TYPE},
    {"name": "event", "type": [APPLICATION_INITED, APPLICATION_FINISHED, TASK_STARTED, TASK_FINISHED]},
    {"name": "timestamp", "type": "long"}]}


def _now_ms() -> int:
    return int(time.time() * 1000)


def _metrics(m: Optional[Dict[str, float]]) -> List[dict]:
    return [{"name": k, "value": float(v)} for k, v in (m or {}).items()]


def application_inited(app_id: str, num_tasks: int, host: str, container_id: Optional[str] = None) -> dict:
    return {"type": "APPLICATION_INITED", "timestamp": _now_ms(), "event": {
        "__type__": "ApplicationInited", "applicationId": app_id, "numTasks": num_tasks, "host": host,
        "containerID": container_id}}


def application_finished(app_id: str, finished: int, failed: int, metrics: Optional[Dict[str, float]] = None) -> dict:
    return {"type": "APPLICATION_FINISHED", "timestamp": _now_ms(), "event": {
        "__type__": "ApplicationFinished", "applicationId": app_id, "finishedTasks": finished,
        "failedTasks": failed, "metrics": _metrics(metrics)}}


def task_started(task_type: str, index: int, host: str, container_id: Optional[str] = None) -> dict:
    return {"type": "TASK_STARTED", "timestamp": _now_ms(), "event": {
        "__type__": "TaskStarted", "taskType": task_type, "taskIndex": index, "host": host,
        "containerID": container_id}}


def task_finished(task_type: str, index: int, status: str, metrics: Optional[Dict[str, float]] = None,
                  diagnostic: Optional[str] = None) -> dict:
    return {"type": "TASK_FINISHED", "timestamp": _now_ms(), "event": {
        "__type__": "TaskFinished", "taskType": task_type, "taskIndex": index, "status": status,
        "metrics": _metrics(metrics), "containerDiagnostic": diagnostic}}
