"""Cluster abstraction (reference: ClusterInterface/Interfaces.cs:31-330 — ICluster, IScheduler,
IComputer, IProcess, IProcessWatcher, Affinity — and ClusterInterface/HttpScheduler.cs:34-126).

On one 8xB200 box the "computers" are GPUs (one worker process each) and the single "rack" is the
NVSwitch domain; `LocalCluster` implements the interface over the native delay scheduler."""
from __future__ import annotations

import abc
import enum
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

from ..runtime.services import LocalScheduler


class ProcessState(enum.Enum):
    QUEUED, MATCHED, CREATED, STARTED, EXITED = range(5)


class ExitState(enum.Enum):
    SCHEDULE_FAILED, STATUS_FAILED, SCHEDULE_CANCELED, PROCESS_EXITED = range(4)


@dataclass
class Computer:
    name: str
    index: int
    rack: str = "nvswitch0"
    host: str = "localhost"
    directory: str = ""


@dataclass
class Affinity:
    """isHardConstraint + localities + weight (Interfaces.cs Affinity)."""
    computers: List[int] = field(default_factory=list)
    weight: int = 0
    hard: bool = False


class IProcessWatcher(abc.ABC):
    def on_queued(self, proc): ...
    def on_matched(self, proc, computer: Computer): ...
    def on_created(self, proc): ...
    def on_started(self, proc): ...
    def on_exited(self, proc, state: ExitState, exit_code: int, error: str): ...


@dataclass
class Process:
    id: int
    command_line: str
    affinity: Affinity
    watcher: Optional[IProcessWatcher] = None
    state: ProcessState = ProcessState.QUEUED
    computer: Optional[Computer] = None


class ICluster(abc.ABC):
    @abc.abstractmethod
    def get_computers(self) -> List[Computer]: ...
    @abc.abstractmethod
    def schedule_process(self, process: Process) -> None: ...
    @abc.abstractmethod
    def cancel_process(self, process: Process) -> None: ...
    @abc.abstractmethod
    def local_path(self, computer: Computer, name: str, compression: str = "None") -> str: ...


class LocalCluster(ICluster):
    """GPUs of one box as computers; delay scheduling decides which idle GPU takes which process."""

    def __init__(self, ngpus: int, work_dir: str = "/tmp", rack_delay: float = 1.0, cluster_delay: float = 2.0):
        self.computers = [Computer(f"gpu{i}", i, directory=work_dir) for i in range(ngpus)]
        self.scheduler = LocalScheduler([0] * ngpus, rack_delay, cluster_delay)
        self._procs: Dict[int, Process] = {}

    def get_computers(self):
        return list(self.computers)

    def schedule_process(self, process: Process) -> None:
        self._procs[process.id] = process
        self.scheduler.submit(process.id, process.affinity.computers, process.affinity.hard)
        if process.watcher:
            process.watcher.on_queued(process)

    def cancel_process(self, process: Process) -> None:
        p = self._procs.pop(process.id, None)
        if p is not None and p.watcher:
            p.watcher.on_exited(p, ExitState.SCHEDULE_CANCELED, -1, "cancelled")

    def computer_idle(self, index: int, age_bias: float = 0.0) -> Optional[Process]:
        """An idle GPU claims the best waiting process (two-sided matching)."""
        pid = self.scheduler.computer_idle(index, age_bias)
        if pid is None or pid not in self._procs:
            return None
        p = self._procs[pid]
        p.computer, p.state = self.computers[index], ProcessState.MATCHED
        if p.watcher:
            p.watcher.on_matched(p, p.computer)
        return p

    def local_path(self, computer: Computer, name: str, compression: str = "None") -> str:
        return f"file://{computer.directory}/{name}?c={compression}"

    def remote_path(self, computer: Computer, name: str, file_server: str) -> str:
        """Where ANOTHER computer reads `name` from: through the producing node's file server
        (ClusterInterface/HttpScheduler.cs:65-91); `file_server` is that node's process-service URI."""
        return f"{file_server.rstrip('/')}/file/{name.lstrip('/')}"
