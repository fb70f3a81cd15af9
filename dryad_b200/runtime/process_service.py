"""Node daemon: process table, versioned key/value mailbox and file server over HTTP.

Reference: ProcessService/ProcessService.cs:52-752 (Launch, BlockOnStatus, SetValue, Kill),
ProcessService/HttpServer.cs:42-707 (PUT create/launch/setvalue/kill, GET long-poll
`?key&version&timeout`, ranged file GET), headers X-Dryad-ValueVersion / X-Dryad-ProcessStatus
(ClusterInterface/Process.cs:236-240).  The mailbox itself is the native one (csrc/host/runtime.cpp);
this module is its HTTP face so a worker on another process can long-poll its command.
"""
from __future__ import annotations

import os
import subprocess
import threading
import urllib.parse
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Dict, Optional

from .services import Mailbox

H_VERSION, H_STATUS, H_EXIT = "X-Dryad-ValueVersion", "X-Dryad-ProcessStatus", "X-Dryad-ProcessExitCode"


class ProcessService:
    def __init__(self, root_dir: str, host: str = "127.0.0.1", port: int = 0):
        self.root = os.path.abspath(root_dir)
        os.makedirs(self.root, exist_ok=True)
        self.mailboxes: Dict[str, Mailbox] = {}
        self.procs: Dict[str, subprocess.Popen] = {}
        self._lock = threading.Lock()
        svc = self

        class Handler(BaseHTTPRequestHandler):
            protocol_version = "HTTP/1.1"       # keep-alive: a worker keeps ONE connection for all its polls and reports

            def log_message(self, *a):
                pass

            def setup(self):
                super().setup()
                import socket
                self.request.setsockopt(socket.IPPROTO_TCP, socket.TCP_NODELAY, 1)   # small messages: no 40 ms delayed-ACK stalls

            def end_headers(self):
                if not any(h.lower().startswith(b"content-length") for h in getattr(self, "_headers_buffer", [])):
                    self.send_header("Content-Length", "0")
                super().end_headers()

            def _mb(self, pid: str) -> Mailbox:
                with svc._lock:
                    m = svc.mailboxes.get(pid)
                    if m is None:
                        m = svc.mailboxes[pid] = Mailbox()
                    return m

            def do_PUT(self):
                u = urllib.parse.urlparse(self.path)
                q = urllib.parse.parse_qs(u.query)
                parts = u.path.strip("/").split("/")
                body = self.rfile.read(int(self.headers.get("Content-Length", 0) or 0))
                if len(parts) == 3 and parts[0] == "process" and parts[2] == "setvalue":
                    mb = self._mb(parts[1])
                    v = mb.set(q["key"][0], q.get("status", [""])[0], body)
                    got = None
                    if "wait" in q:
                        # report-and-wait: the same exchange long-polls another key (a worker hands in its
                        # status and receives its next command in one round trip)
                        got = mb.block_on_status(q["wait"][0], int(q.get("version", ["0"])[0]), float(q.get("timeout", ["1"])[0]))
                    self.send_response(200)
                    self.send_header(H_VERSION, str(got[0] if got else v))
                    if got:
                        self.send_header(H_STATUS, got[1])
                        self.send_header("X-Dryad-Waited", "1")
                        self.send_header("Content-Length", str(len(got[2])))
                    self.end_headers()
                    if got:
                        self.wfile.write(got[2])
                elif len(parts) == 3 and parts[0] == "process" and parts[2] == "launch":
                    cmd = body.decode().split("\x00")
                    log = open(os.path.join(svc.root, f"process-{parts[1]}-stdout.txt"), "wb")
                    with svc._lock:
                        svc.procs[parts[1]] = subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT, cwd=svc.root)
                    self.send_response(200)
                    self.end_headers()
                elif len(parts) == 3 and parts[0] == "process" and parts[2] == "kill":
                    with svc._lock:
                        p = svc.procs.get(parts[1])
                    if p is not None and p.poll() is None:
                        p.kill()
                    self.send_response(200)
                    self.end_headers()
                else:
                    self.send_response(404)
                    self.end_headers()

            def do_GET(self):
                u = urllib.parse.urlparse(self.path)
                q = urllib.parse.parse_qs(u.query)
                parts = u.path.strip("/").split("/")
                if len(parts) == 3 and parts[0] == "process" and parts[2] == "status":
                    got = self._mb(parts[1]).block_on_status(q["key"][0], int(q.get("version", ["0"])[0]),
                                                             float(q.get("timeout", ["1"])[0]))
                    with svc._lock:
                        p = svc.procs.get(parts[1])
                    self.send_response(200 if got else 204)
                    if got:
                        self.send_header(H_VERSION, str(got[0]))
                        self.send_header(H_STATUS, got[1])
                    if p is not None and p.poll() is not None:
                        self.send_header(H_EXIT, str(p.returncode))
                    if got:
                        self.send_header("Content-Length", str(len(got[2])))
                    self.end_headers()
                    if got:
                        self.wfile.write(got[2])
                elif parts and parts[0] == "file":
                    path = os.path.normpath(os.path.join(svc.root, *parts[1:]))
                    if not path.startswith(svc.root) or not os.path.isfile(path):
                        self.send_response(404)
                        self.end_headers()
                        return
                    size = os.path.getsize(path)
                    start, end = 0, size - 1
                    rng = self.headers.get("Range")
                    if rng and rng.startswith("bytes="):
                        a, _, b = rng[6:].partition("-")
                        start = int(a or 0)
                        end = int(b) if b else size - 1
                    with open(path, "rb") as f:
                        f.seek(start)
                        data = f.read(end - start + 1)
                    self.send_response(206 if rng else 200)
                    self.send_header("Content-Length", str(len(data)))
                    self.end_headers()
                    self.wfile.write(data)
                else:
                    self.send_response(404)
                    self.end_headers()

        class Server(ThreadingHTTPServer):
            # every worker of a job long-polls and reports at the same instant: the default listen
            # backlog of 5 drops connections (1 s SYN retransmits) from 8 ranks on
            request_queue_size = 256
            daemon_threads = True

        self.httpd = Server((host, port), Handler)
        self.port = self.httpd.server_address[1]
        self.uri = f"http://{host}:{self.port}"
        self._thread = threading.Thread(target=self.httpd.serve_forever, daemon=True)
        self._thread.start()

    def close(self):
        self.httpd.shutdown()
        self.httpd.server_close()
        for p in self.procs.values():
            if p.poll() is None:
                p.kill()
        for m in self.mailboxes.values():
            m.close()
