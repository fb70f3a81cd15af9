"""Leveled logging (reference: LinqToDryad/DryadLinqLog.cs:30-190 — vertex log
`vertex-<id>-<ver>-LinqLog.txt` with levels Critical..Verbose, client log `DryadLinqClient.log`;
level from DRYAD_LOGGING_LEVEL, LinqToDryad/Constants.cs:72-85)."""
from __future__ import annotations

import datetime
import os
import threading

LEVELS = {"OFF": 0, "CRITICAL": 1, "ERROR": 2, "WARN": 3, "INFO": 4, "VERBOSE": 5}


class DryadLinqLog:
    def __init__(self, path: str, level: str = None):
        self.path = path
        self.level = LEVELS.get((level or os.environ.get("DRYAD_LOGGING_LEVEL", "ERROR")).upper(), 2)
        self._lock = threading.Lock()
        self._f = None

    def _write(self, lvl: str, msg: str, *args) -> None:
        if LEVELS[lvl] > self.level:
            return
        with self._lock:
            if self._f is None:
                os.makedirs(os.path.dirname(os.path.abspath(self.path)) or ".", exist_ok=True)
                self._f = open(self.path, "a")
            ts = datetime.datetime.now(datetime.timezone.utc).strftime("%Y-%m-%d %H:%M:%S.%f")
            self._f.write(f"{ts} {lvl[0]} {msg.format(*args) if args else msg}\n")
            self._f.flush()

    def critical(self, m, *a): self._write("CRITICAL", m, *a)
    def error(self, m, *a): self._write("ERROR", m, *a)
    def warning(self, m, *a): self._write("WARN", m, *a)
    def info(self, m, *a): self._write("INFO", m, *a)
    def verbose(self, m, *a): self._write("VERBOSE", m, *a)

    @staticmethod
    def for_vertex(log_dir: str, vertex_id: int, version: int, level: str = None) -> "DryadLinqLog":
        return DryadLinqLog(os.path.join(log_dir, f"vertex-{vertex_id}-{version}-LinqLog.txt"), level)

    @staticmethod
    def client(log_dir: str, level: str = None) -> "DryadLinqLog":
        return DryadLinqLog(os.path.join(log_dir, "DryadLinqClient.log"), level)
