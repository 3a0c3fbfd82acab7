"""History portal: a small HTTP server over the job-history tree.

Routes kept from tony-portal (conf/routes:1-5): ``/`` jobs table, ``/config/:jobId``,
``/jobs/:jobId`` (events), ``/logs/:jobId``; plus ``/api/...`` JSON twins.  Background duties
kept too: warm metadata/config/event/log caches keyed by ``application_\\d+_\\d+``
(app/cache/CacheWrapper.java:30,69-127), move finished jobs to ``finished/yyyy/MM/dd`` and
mark killed ones (app/history/HistoryFileMover.java:74-168), purge past retention
(app/history/HistoryFilePurger.java:49-112), create the history directories at start-up
(app/hadoop/Requirements.java:83-106).  Play/Twirl/Guava become http.server + an LRU dict.

    python -m tony_b200.portal.server --history /path/to/tony-history [--port 19886]
"""
from __future__ import annotations

import argparse
import html
import json
import os
import re
import threading
import time
from collections import OrderedDict
from dataclasses import asdict
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Callable, List, Optional

from ..conf import keys as K
from ..conf.configuration import TonyConfiguration
from ..events import history

JOB_ID_RE = re.compile(r"^application_\d+_\d+$")


class LruCache:
    def __init__(self, max_entries: int):
        self.max = max_entries
        self._d: "OrderedDict[str, object]" = OrderedDict()
        self._lock = threading.Lock()

    def get(self, k):
        with self._lock:
            if k in self._d:
                self._d.move_to_end(k)
                return self._d[k]
            return None

    def put(self, k, v):
        with self._lock:
            self._d[k] = v
            self._d.move_to_end(k)
            while len(self._d) > self.max:
                self._d.popitem(last=False)

    def values(self):
        with self._lock:
            return list(self._d.values())


class Portal:
    def __init__(self, conf: TonyConfiguration, killed_apps: Callable[[], set] = lambda: set()):
        self.conf = conf
        root = conf.get(K.HISTORY_LOCATION)
        self.intermediate = conf.get(K.HISTORY_INTERMEDIATE) or os.path.join(root, "intermediate")
        self.finished = conf.get(K.HISTORY_FINISHED) or os.path.join(root, "finished")
        for d in (self.intermediate, self.finished):
            os.makedirs(d, exist_ok=True)
        n = conf.get_int(K.PORTAL_CACHE_MAX_ENTRIES, 10000)
        self.metadata, self.configs, self.events, self.logs = (LruCache(n) for _ in range(4))
        self.mover = history.HistoryFileMover(self.intermediate, self.finished,
                                              conf.get(K.HISTORY_FINISHED_DIR_TIMEZONE, "UTC"),
                                              killed_apps, self.update_caches)
        self.purger = history.HistoryFilePurger(self.finished, conf.get_int(K.HISTORY_RETENTION_SEC, 2592000))
        self._stop = threading.Event()
        self._threads: List[threading.Thread] = []
        self.httpd: Optional[ThreadingHTTPServer] = None

    # -- caches -------------------------------------------------------------------------
    def update_caches(self, job_dir: str) -> None:
        job_id = os.path.basename(job_dir.rstrip("/"))
        if not JOB_ID_RE.match(job_id):
            return
        md = history.parse_metadata(job_dir)
        if md is None:
            return
        self.metadata.put(job_id, md)
        self.configs.put(job_id, history.parse_config(job_dir))
        self.events.put(job_id, history.parse_events(job_dir))
        self.logs.put(job_id, history.parse_logs(job_dir, md.user))

    def warm_up(self) -> None:
        for root in (self.finished, self.intermediate):
            for dirpath, dirs, _ in os.walk(root):
                for d in dirs:
                    if JOB_ID_RE.match(d):
                        self.update_caches(os.path.join(dirpath, d))

    def _lookup(self, cache: LruCache, job_id: str):
        if not JOB_ID_RE.match(job_id):
            return None
        hit = cache.get(job_id)
        if hit is not None:
            return hit
        job_dir = history.find_job_dir(self.finished, self.intermediate, job_id)
        if job_dir is None:
            return None
        self.update_caches(job_dir)
        return cache.get(job_id)

    def jobs(self) -> List[history.JobMetadata]:
        return sorted(self.metadata.values(), key=lambda m: (-m.completed, -m.started, m.user))

    # -- background threads ---------------------------------------------------------------
    def _every(self, interval_s: float, fn: Callable[[], object], name: str) -> None:
        def loop():
            while not self._stop.wait(interval_s):
                try:
                    fn()
                except Exception:  # noqa: BLE001
                    pass

        t = threading.Thread(target=loop, name=name, daemon=True)
        t.start()
        self._threads.append(t)

    def start(self, port: int = 0, host: str = "127.0.0.1") -> int:
        self.warm_up()
        self.mover.run_once()
        self._every(self.conf.get_int(K.HISTORY_MOVER_INTERVAL_MS, 300000) / 1000.0, self.mover.run_once, "mover")
        self._every(self.conf.get_int(K.HISTORY_PURGER_INTERVAL_MS, 21600000) / 1000.0, self.purger.run_once, "purger")
        portal = self

        class Handler(BaseHTTPRequestHandler):
            def log_message(self, *a):
                pass

            def _send(self, code: int, body: str, ctype: str = "text/html; charset=utf-8"):
                data = body.encode("utf-8")
                self.send_response(code)
                self.send_header("Content-Type", ctype)
                self.send_header("Content-Length", str(len(data)))
                self.end_headers()
                self.wfile.write(data)

            def do_GET(self):
                code, body, ctype = portal.route(self.path)
                self._send(code, body, ctype)

        self.httpd = ThreadingHTTPServer((host, port), Handler)
        self.httpd.daemon_threads = True
        threading.Thread(target=self.httpd.serve_forever, name="portal-http", daemon=True).start()
        return self.httpd.server_address[1]

    def stop(self) -> None:
        self._stop.set()
        if self.httpd:
            self.httpd.shutdown()
            self.httpd.server_close()

    # -- rendering --------------------------------------------------------------------------
    @staticmethod
    def _page(title: str, head: List[str], rows: List[List[str]]) -> str:
        th = "".join(f"<th>{html.escape(h)}</th>" for h in head)
        trs = "".join("<tr>" + "".join(f"<td>{c}</td>" for c in r) + "</tr>" for r in rows)
        return (f"<!doctype html><html><head><title>{html.escape(title)}</title>"
                "<style>body{font-family:sans-serif}table{border-collapse:collapse}"
                "td,th{border:1px solid #ccc;padding:4px 8px}</style></head>"
                f"<body><h2>{html.escape(title)}</h2><p><a href='/'>all jobs</a></p>"
                f"<table><thead><tr>{th}</tr></thead><tbody>{trs}</tbody></table></body></html>")

    @staticmethod
    def _ts(ms: int) -> str:
        return "-" if ms < 0 else time.strftime("%d %b %Y %H:%M:%S", time.gmtime(ms / 1000.0))

    def route(self, path: str):
        path = path.split("?", 1)[0].rstrip("/") or "/"
        api = path.startswith("/api")
        if api:
            path = path[4:] or "/"
        e = html.escape
        if path == "/":
            jobs = self.jobs()
            if api:
                return 200, json.dumps([asdict(j) for j in jobs]), "application/json"
            rows = [[f"<a href='{e(j.job_link)}'>{e(j.id)}</a>", f"<a href='{e(j.config_link)}'>config</a>",
                     f"<a href='/logs/{e(j.id)}'>logs</a>", self._ts(j.started), self._ts(j.completed),
                     e(j.user), e(j.status)] for j in jobs]
            return 200, self._page("TonY jobs", ["Job", "Config", "Logs", "Started", "Completed", "User", "Status"],
                                   rows), "text/html; charset=utf-8"
        m = re.match(r"^/(config|jobs|logs)/([^/]+)$", path)
        if not m:
            return 404, "not found", "text/plain"
        kind, job_id = m.groups()
        cache = {"config": self.configs, "jobs": self.events, "logs": self.logs}[kind]
        data = self._lookup(cache, job_id)
        if data is None:
            return 404, f"no history for {e(job_id)}", "text/plain"
        if api:
            return 200, json.dumps([asdict(x) for x in data]), "application/json"
        if kind == "config":
            rows = [[e(c.name), e(c.value), str(c.final).lower(), e(c.source)] for c in data]
            return 200, self._page(f"Configuration of {job_id}", ["Name", "Value", "Final", "Source"], rows), "text/html; charset=utf-8"  # noqa
        if kind == "jobs":
            rows = [[e(ev.type), e(json.dumps({k: v for k, v in ev.event.items() if k != '__type__'})),
                     self._ts(ev.timestamp)] for ev in data]
            return 200, self._page(f"Events of {job_id}", ["Type", "Event", "Timestamp"], rows), "text/html; charset=utf-8"
        rows = [[e(l.task_type), str(l.task_index), e(l.host), e(l.container_id),
                 f"<a href='{e(l.url)}'>{e(l.url)}</a>"] for l in data]
        return 200, self._page(f"Logs of {job_id}", ["Task", "Index", "Host", "Container", "Log"], rows), "text/html; charset=utf-8"  # noqa


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="TonY history portal")
    ap.add_argument("--history", help="history root (overrides tony.history.location)")
    ap.add_argument("--conf_file", help="tony-site.xml style configuration")
    ap.add_argument("--port", type=int, default=None)
    ap.add_argument("--host", default="127.0.0.1")
    args = ap.parse_args(argv)
    conf = TonyConfiguration()
    if args.conf_file:
        conf.add_resource(args.conf_file)
    if args.history:
        conf.set(K.HISTORY_LOCATION, args.history)
        conf.set(K.HISTORY_INTERMEDIATE, os.path.join(args.history, "intermediate"))
        conf.set(K.HISTORY_FINISHED, os.path.join(args.history, "finished"))
    port = args.port
    if port is None:
        hp = conf.get(K.HTTP_PORT, "disabled")
        port = int(hp) if hp and hp.isdigit() else conf.get_int(K.HTTPS_PORT, 19886)
    portal = Portal(conf)
    actual = portal.start(port, args.host)
    print(f"tony portal listening on http://{args.host}:{actual}", flush=True)
    try:
        while True:
            time.sleep(3600)
    except KeyboardInterrupt:
        portal.stop()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
