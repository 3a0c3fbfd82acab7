#!/usr/bin/env python
"""Horovod rendezvous driver process (run by the ``driver`` task).

Contract kept from the reference's helper (tony-core/src/main/resources/horovod_driver.py
:26-83,124-188 and horovod/HorovodDriver.java:56,128-216): given ``-w host1:n1,host2:n2`` it
starts a rendezvous server, computes one slot per worker process
(rank / localRank / crossRank / size / localSize / crossSize), and publishes them by
creating ``<dir>/<port>____HOROVOD_RENDEZVOUS_SERVER____`` holding the slot list as JSON, then
stays alive until killed.  ``-t -p PORT`` serves a fixed fake 2-slot plan without touching the
network, ``-f`` exits non-zero at once (fault-injection).

Re-designed rather than wrapped: the slot plan is computed here, and when the ``horovod``
package is absent (it is not installed on the B200 image) a small built-in HTTP key-value
store speaks the PUT/GET/DELETE protocol Horovod's Gloo controller expects, so the
``HOROVOD_GLOO_RENDEZVOUS_*`` contract still points at a live server.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer

PORT_FILE_SUFFIX = "____HOROVOD_RENDEZVOUS_SERVER____"


def parse_worker_list(spec: str):
    hosts = []
    for item in spec.split(","):
        item = item.strip()
        if not item:
            continue
        host, _, n = item.rpartition(":")
        hosts.append((host, int(n)))
    return hosts


def host_assignments(hosts):
    """Ranks are dealt host by host; crossRank = position of the host among hosts that have
    that local rank (same plan shape as horovod.runner.common.util.hosts.get_host_assignments)."""
    size = sum(n for _, n in hosts)
    slots = []
    rank = 0
    for hi, (host, n) in enumerate(hosts):
        for local in range(n):
            cross_hosts = [h for h, m in hosts if m > local]
            slots.append({
                "hostname": host, "rank": rank, "localRank": local,
                "crossRank": cross_hosts.index(host) if host in cross_hosts else 0,
                "size": size, "localSize": n, "crossSize": len(cross_hosts),
            })
            rank += 1
    return slots


def fake_plan(first_host: str):
    return [
        {"hostname": first_host, "rank": 0, "localRank": 0, "crossRank": 0, "size": 2, "localSize": 2, "crossSize": 1},
        {"hostname": first_host, "rank": 1, "localRank": 1, "crossRank": 0, "size": 2, "localSize": 2, "crossSize": 1},
    ]


class _KVHandler(BaseHTTPRequestHandler):
    store = {}
    lock = threading.Lock()

    def log_message(self, *a):  # quiet
        pass

    def do_GET(self):
        with self.lock:
            v = self.store.get(self.path)
        if v is None:
            self.send_response(404)
            self.end_headers()
            return
        self.send_response(200)
        self.send_header("Content-Length", str(len(v)))
        self.end_headers()
        self.wfile.write(v)

    def do_PUT(self):
        n = int(self.headers.get("Content-Length", "0"))
        body = self.rfile.read(n)
        with self.lock:
            self.store[self.path] = body
        self.send_response(200)
        self.end_headers()

    do_POST = do_PUT

    def do_DELETE(self):
        with self.lock:
            self.store.pop(self.path, None)
        self.send_response(200)
        self.end_headers()


def start_rendezvous(slots) -> int:
    """Return the port of a live rendezvous server initialised with ``slots``."""
    try:  # real Horovod, if the user's environment has it
        from horovod.runner.common.util.hosts import SlotInfo  # type: ignore
        from horovod.runner.http.http_server import RendezvousServer  # type: ignore

        server = RendezvousServer(verbose=1)
        port = server.start()
        plan = [SlotInfo(hostname=s["hostname"], rank=s["rank"], local_rank=s["localRank"],
                         cross_rank=s["crossRank"], size=s["size"], local_size=s["localSize"],
                         cross_size=s["crossSize"]) for s in slots]
        server.init(plan)
        return port
    except Exception:  # noqa: BLE001 - fall through to the built-in store
        pass
    httpd = ThreadingHTTPServer(("0.0.0.0", 0), _KVHandler)
    httpd.daemon_threads = True
    threading.Thread(target=httpd.serve_forever, daemon=True).start()
    # publish the plan the way workers look it up: /rendezvous/rank_and_size/<host>:<local_rank>
    for s in slots:
        key = f"/rendezvous/rank_and_size/{s['hostname']}:{s['localRank']}"
        _KVHandler.store[key] = json.dumps(s).encode()
    return httpd.server_address[1]


def write_port_file(out_dir: str, port: int, slots) -> str:
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, f"{port}{PORT_FILE_SUFFIX}")
    tmp = path + ".tmp"
    with open(tmp, "w") as f:
        json.dump(slots, f)
    os.replace(tmp, path)
    return path


def main(argv=None) -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("-w", "--worker_list", required=True)
    ap.add_argument("-t", "--test_mode", action="store_true")
    ap.add_argument("-p", "--fake_port", type=int, default=9999)
    ap.add_argument("-f", "--fail_in_test", action="store_true")
    ap.add_argument("-o", "--output_dir", default=os.environ.get("DRIVER_OUTPUT_PATH") or os.getcwd())
    args = ap.parse_args(argv)
    hosts = parse_worker_list(args.worker_list)
    if args.fail_in_test:
        sys.stderr.write("horovod driver: failing on request (test mode)\n")
        return 1
    if args.test_mode:
        slots, port = fake_plan(hosts[0][0] if hosts else "localhost"), args.fake_port
    else:
        slots = host_assignments(hosts)
        port = start_rendezvous(slots)
    path = write_port_file(args.output_dir, port, slots)
    print(f"horovod driver ready: port={port} slots={len(slots)} file={path}", flush=True)
    try:
        while True:
            time.sleep(3600)
    except KeyboardInterrupt:
        return 0
    finally:
        try:
            os.unlink(path)
        except OSError:
            pass


if __name__ == "__main__":
    raise SystemExit(main())
