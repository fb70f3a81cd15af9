"""Control plane of an SPMD job: rank 0 is the graph manager, every rank is a worker.

Roles of the reference kept here (nothing of its code): the GM decides which stage attempt runs next
and with which VERSION (GraphManager/vertex/DrVertex.cpp:1042-1171, DrGraph.cpp:392-456); a start
command reaches a worker through the node daemon's versioned key/value mailbox, which the worker
long-polls (ProcessService/ProcessService.cs:674-746, DryadVertex/VertexHost/system/dprocess/src/
dvertexpncontrol.h:34-106); the worker answers with a status block
(GraphManager/vertex/DrVertexRecord.cpp:849-871).

  rank 0   native GraphManager (csrc/host/sched.cpp) with one GM vertex per plan stage — the `world`
           partitions of an SPMD stage run collectives together, so they are one gang: they start
           together and a failure of any member re-runs the whole stage with a new version —,
           a ProcessService (HTTP face of the native mailbox) and the decision loop.
  rank r   long-polls mailbox `rank-r`, key "cmd", for the next TLV launch descriptor
           (runtime/command.py), runs that stage attempt, PUTs its status block to key "status".

Inside a stage attempt, `agree()` is the only place where ranks learn about each other's failures:
it is called before every collective (host or device) and once at the end of the attempt.  A rank
whose vertex code raised calls it once with ok=False from its handler; it pairs with whatever
agreement point the healthy ranks reach next, so nobody is left waiting in a collective the failed
rank will never enter.
"""
from __future__ import annotations

import time
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist

from ..api.errors import DryadLinqException, ErrorCode
from .command import CMD_START, CMD_TERMINATE, STATE_COMPLETED, STATE_FAILED, VertexCommand, terminate_command  # noqa: F401

class StageAborted(Exception):
    """Another rank failed in this stage attempt: unwind, the GM decides what happens next."""


class RankLost(DryadLinqException):
    def __init__(self, message: str):
        super().__init__(ErrorCode.DidNotCompleteSuccessfully, message)


class ControlPlane:
    """One per context (the HTTP service and its port live as long as the context)."""

    def __init__(self, context, device):
        self.rank, self.world = context.rank, context.world
        self.device = device
        self.timeout = float(getattr(context, "spmd_control_timeout", 120.0))
        self.service = None
        # Commands are a SEQUENCE, one mailbox key per command ("cmd/1", "cmd/2", ...).  A mailbox key holds only
        # its latest value, so a single "cmd" key loses a command when the manager posts two before a rank has
        # polled — which it does at job boundaries (Terminate of job k, then the first Start of job k+1): the
        # rank then ran job k+1's stages against job k's plan.  Every rank receives every command, so both
        # sides count in lockstep.
        self._posted = 0
        self._received = 0
        self._status_seen = [0] * self.world
        uri: List[Optional[str]] = [None]
        if self.rank == 0:
            import tempfile
            from .process_service import ProcessService
            self.service = ProcessService(tempfile.mkdtemp(prefix="dryad_ps_"))
            uri[0] = self.service.uri
        dist.broadcast_object_list(uri, src=0)            # bootstrap only: where the mailbox lives
        self.uri = uri[0]
        self.commands_delivered = 0
        self._conn = None                     # persistent HTTP connection of a worker rank
        self._prefetched = None               # next command, when it arrived with the last status report

    # ---- mailbox access: rank 0 talks to the native mailboxes directly, the others over HTTP ------
    def _mbox(self, r: int):
        from .services import Mailbox
        with self.service._lock:
            m = self.service.mailboxes.get(f"rank-{r}")
            if m is None:
                m = self.service.mailboxes[f"rank-{r}"] = Mailbox()
            return m

    def _client(self):
        if self._conn is None:
            from .process_service import MailboxClient
            self._conn = MailboxClient(self.uri, f"rank {self.rank}")
        return self._conn

    def post_command(self, cmd: VertexCommand) -> None:
        """GM side: deliver one launch descriptor to every rank's mailbox."""
        data = cmd.serialize()
        self._posted += 1
        for r in range(self.world):
            self._mbox(r).set(f"cmd/{self._posted}", "start" if cmd.command == CMD_START else "terminate", data)
        self.commands_delivered += 1

    def next_command(self) -> VertexCommand:
        """Worker side: long-poll this rank's mailbox for a command newer than the last one seen."""
        deadline = time.monotonic() + self.timeout
        while True:
            left = deadline - time.monotonic()
            if left <= 0:
                raise RankLost(f"rank {self.rank}: no command from the graph manager for {self.timeout:.0f} s")
            poll = min(left, 5.0)
            key = f"cmd/{self._received + 1}"
            if self._prefetched is not None:                  # delivered together with the last status report
                got, self._prefetched = self._prefetched, None
            elif self.rank == 0:
                got = self._mbox(0).block_on_status(key, 0, poll)
            else:
                got = self._http_get(f"rank-{self.rank}", key, 0, poll)
            if got:
                self._received += 1
                cmd = VertexCommand.parse(got[2])
                if cmd is None:
                    raise DryadLinqException(ErrorCode.Internal, "malformed launch descriptor")
                return cmd

    def report(self, uid: int, version: int, ok: bool, message: str = "") -> None:
        st = VertexCommand(CMD_START, uid, version, state=STATE_COMPLETED if ok else STATE_FAILED,
                           exit_code=0 if ok else 1, message=message[:4000])
        data = st.serialize()
        if self.rank == 0:
            self._mbox(0).set("status", "completed" if ok else "failed", data)
        else:
            # report and long-poll the next command in ONE exchange
            try:
                self._prefetched = self._client().set(f"rank-{self.rank}", "status", "completed" if ok else "failed", data,
                                                      wait=f"cmd/{self._received + 1}", seen=0, wait_timeout=5.0, timeout=self.timeout)
            except ConnectionError as ex:
                raise RankLost(str(ex)) from ex

    def collect(self, uid: int, version: int) -> Tuple[bool, str]:
        """GM side: wait for every rank's status block of (stage, version).  A rank that does not
        answer within the timeout is reported lost."""
        ok, first_error = True, ""
        for r in range(self.world):
            deadline = time.monotonic() + self.timeout
            while True:
                left = deadline - time.monotonic()
                if left <= 0:
                    return False, f"rank {r} did not report stage {uid} v{version} within {self.timeout:.0f} s (rank lost)"
                got = self._mbox(r).block_on_status("status", self._status_seen[r], min(left, 5.0))
                if not got:
                    continue
                self._status_seen[r] = got[0]
                st = VertexCommand.parse(got[2])
                if st is None or st.vertex_id != uid or st.version != version:
                    continue                                  # a stale report of an earlier attempt
                if st.state != STATE_COMPLETED:
                    ok = False
                    if st.message and not st.message.startswith("aborted:") and not first_error:
                        first_error = f"rank {r}: {st.message}"
                break
        return ok, first_error

    def _http_get(self, pid: str, key: str, seen: int, timeout: float):
        try:
            return self._client().get(pid, key, seen, timeout)
        except ConnectionError as ex:
            raise RankLost(str(ex)) from ex

    # ---- agreement point ------------------------------------------------------------------------
    def agree(self, ok: bool = True) -> bool:
        """All-reduce(MIN) of one flag: True when every rank is still healthy in this attempt."""
        t = torch.tensor([1 if ok else 0], dtype=torch.int64, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(int(t.item()))

    def close(self) -> None:
        if self._conn is not None:
            self._conn.close()
            self._conn = None
        if self.service is not None:
            self.service.close()
            self.service = None
