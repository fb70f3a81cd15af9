"""Job submission back ends (reference: LinqToDryad/IDryadLinqJobSubmission.cs:30-54,
DryadLinqJobExecutor.cs:36-215, LocalJobSubmission.cs:97-363).

`LocalJobSubmission` is the local MULTI-PROCESS mode: it creates `<DRYAD_HOME>/LocalJobs/<n>/log`,
launches `torchrun` with one worker process per GPU (or per CPU rank over gloo) running a job script,
passes DRYAD_THREADS_PER_WORKER / DRYAD_LOGGING_LEVEL, and derives the job status from the process
exit code plus `log/error.txt` — the torchrun launcher plays the Peloponnese process manager.
"""
from __future__ import annotations

import abc
import os
import subprocess
import sys
from typing import Dict, List, Optional

from .context import JobStatus


class IDryadLinqJobSubmission(abc.ABC):
    @abc.abstractmethod
    def add_job_option(self, name: str, value: str) -> None: ...
    @abc.abstractmethod
    def add_local_file(self, path: str) -> None: ...
    @abc.abstractmethod
    def submit_job(self) -> None: ...
    @abc.abstractmethod
    def get_status(self) -> JobStatus: ...
    @abc.abstractmethod
    def terminate_job(self) -> None: ...
    @abc.abstractmethod
    def wait_for_completion(self, timeout: Optional[float] = None) -> JobStatus: ...
    @abc.abstractmethod
    def get_job_id(self) -> str: ...


class LocalJobSubmission(IDryadLinqJobSubmission):
    _counter = 0

    def __init__(self, script: str, args: Optional[List[str]] = None, num_processes: int = 2, backend: str = "gloo",
                 dryad_home: Optional[str] = None, threads_per_worker: int = 0, logging_level: str = "ERROR",
                 master_port: int = 29750):
        self.script, self.args = script, list(args or [])
        self.n, self.backend = num_processes, backend
        self.home = dryad_home or os.environ.get("DRYAD_HOME") or os.path.join("/tmp", "DRYADLINQ")
        LocalJobSubmission._counter += 1
        self.job_dir = os.path.join(self.home, "LocalJobs", f"{os.getpid()}_{LocalJobSubmission._counter}")
        self.options: Dict[str, str] = {}
        self.files: List[str] = []
        self.env = {"DRYAD_THREADS_PER_WORKER": str(threads_per_worker or (os.cpu_count() or 1)),
                    "DRYAD_LOGGING_LEVEL": logging_level, "DRYAD_HOME": self.home, "DRYAD_JOB_DIR": self.job_dir,
                    "DRYAD_BACKEND": backend}
        self.port = master_port + LocalJobSubmission._counter
        self.proc: Optional[subprocess.Popen] = None
        self.status = JobStatus.NotSubmitted

    def add_job_option(self, name, value):
        self.options[name] = value
        self.env["DRYAD_OPT_" + name.upper()] = str(value)

    def add_local_file(self, path):
        self.files.append(path)

    def get_job_id(self):
        return os.path.basename(self.job_dir)

    def resource_files(self) -> List[str]:
        """The job's file set: the script, files added with add_local_file and every imported module
        that is not part of the interpreter's own library trees (PeloponneseJobSubmission computes the
        GM / worker file sets by dependency listing, LinqToDryad/PeloponneseJobSubmission.cs:36-179)."""
        from ..plan.type_system import loaded_non_system_modules
        return sorted(set([os.path.abspath(self.script)] + [os.path.abspath(f) for f in self.files] + loaded_non_system_modules()))

    def submit_job(self):
        os.makedirs(os.path.join(self.job_dir, "log"), exist_ok=True)
        with open(os.path.join(self.job_dir, "resources.txt"), "w") as f:
            f.write("\n".join(self.resource_files()) + "\n")
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={self.n}",
               "--master-addr", "127.0.0.1", "--master-port", str(self.port), self.script, *self.args]
        out = open(os.path.join(self.job_dir, "log", "stdout.txt"), "wb")
        self.proc = subprocess.Popen(cmd, stdout=out, stderr=subprocess.STDOUT, env={**os.environ, **self.env})
        self.status = JobStatus.Running

    def get_status(self):
        if self.proc is None:
            return self.status
        rc = self.proc.poll()
        if rc is None:
            return JobStatus.Running
        if self.status == JobStatus.Cancelled:
            return self.status
        err = os.path.join(self.job_dir, "log", "error.txt")
        self.status = JobStatus.Success if rc == 0 and not os.path.exists(err) else JobStatus.Failure
        return self.status

    def terminate_job(self):
        if self.proc is not None and self.proc.poll() is None:
            self.proc.terminate()          # the exact process we started — never a pattern kill
            self.status = JobStatus.Cancelled

    def wait_for_completion(self, timeout=None):
        if self.proc is not None:
            try:
                self.proc.wait(timeout)
            except subprocess.TimeoutExpired:
                return JobStatus.Running
        return self.get_status()

    def error_text(self) -> str:
        err = os.path.join(self.job_dir, "log", "error.txt")
        if os.path.exists(err):
            return open(err).read()
        out = os.path.join(self.job_dir, "log", "stdout.txt")
        return open(out, errors="replace").read()[-2000:] if os.path.exists(out) else ""


class ClusterJobSubmission(IDryadLinqJobSubmission):
    """Cluster mode on one box: a graph-manager PROCESS (`python -m dryad_b200.sched.jm`) that starts
    vertex-host processes through its process service and runs the job package in `job_dir`
    (the role of YarnJobSubmission / PeloponneseJobSubmission: LinqToDryad/YarnJobSubmission.cs,
    PeloponneseJobSubmission.cs:36-179 — stage resources, start the job manager, poll its status)."""

    def __init__(self, job_dir: str, workers: int = 0):
        self.job_dir, self.workers = job_dir, workers
        self.options: Dict[str, str] = {}
        self.files: List[str] = []
        self.proc: Optional[subprocess.Popen] = None
        self.status = JobStatus.NotSubmitted

    def add_job_option(self, name, value):
        self.options[name] = str(value)

    def add_local_file(self, path):
        self.files.append(path)

    def get_job_id(self):
        return os.path.basename(self.job_dir.rstrip("/"))

    def submit_job(self):
        os.makedirs(os.path.join(self.job_dir, "log"), exist_ok=True)
        if self.files:
            import shutil
            res = os.path.join(self.job_dir, "resources")
            os.makedirs(res, exist_ok=True)
            for f in self.files:
                shutil.copy(f, res)
        cmd = [sys.executable, "-m", "dryad_b200.sched.jm", self.job_dir]
        if self.workers:
            cmd += ["--workers", str(self.workers)]
        for k, v in self.options.items():                  # graph-manager options: -mai 64, -at 512MB, -nca ...
            cmd += ["-" + k.lstrip("-")] + ([str(v)] if v not in ("", None, True) else [])
        env = dict(os.environ)
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        env["PYTHONPATH"] = root + os.pathsep + env.get("PYTHONPATH", "")
        out = open(os.path.join(self.job_dir, "log", "jm-stdout.txt"), "wb")
        self.proc = subprocess.Popen(cmd, stdout=out, stderr=subprocess.STDOUT, env=env)
        self.status = JobStatus.Running

    def status_record(self) -> dict:
        import json
        p = os.path.join(self.job_dir, "status.json")
        if not os.path.exists(p):
            return {}
        with open(p) as f:
            return json.load(f)

    def get_status(self):
        if self.proc is None or self.status == JobStatus.Cancelled:
            return self.status
        rc = self.proc.poll()
        if rc is None:
            return JobStatus.Running
        rec = self.status_record()
        self.status = JobStatus.Success if rc == 0 and rec.get("state") == "Success" else JobStatus.Failure
        return self.status

    def terminate_job(self):
        if self.proc is not None and self.proc.poll() is None:
            self.proc.terminate()          # SIGTERM to the graph manager we started: it stops its vertex hosts
            try:
                self.proc.wait(10)
            except subprocess.TimeoutExpired:
                self.proc.kill()
            self.status = JobStatus.Cancelled

    def wait_for_completion(self, timeout=None):
        if self.proc is not None:
            try:
                self.proc.wait(timeout)
            except subprocess.TimeoutExpired:
                return JobStatus.Running
        return self.get_status()

    def error_text(self) -> str:
        rec = self.status_record()
        if rec.get("error"):
            return rec["error"]
        out = os.path.join(self.job_dir, "log", "jm-stdout.txt")
        return open(out, errors="replace").read()[-3000:] if os.path.exists(out) else ""

