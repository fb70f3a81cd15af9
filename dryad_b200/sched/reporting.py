"""Job event log in the reference's `calypso.log` vocabulary (GraphManager/reporting/
DrCalypsoReporting.cpp:84-685): one line per event, UTC timestamp then key=value pairs; vertex
transitions created/starting/running/completed/failed with vertex, version, stagename,
partInStage; job start/stop; final topology.  Consumed by `dryad_b200.tools.jobreport`."""
from __future__ import annotations

import datetime
import os
from typing import Optional


class CalypsoReporter:
    def __init__(self, job_dir: str, level: str = "ERROR"):
        self.job_dir = job_dir
        os.makedirs(os.path.join(job_dir, "log"), exist_ok=True)
        self.path = os.path.join(job_dir, "calypso.log")
        self._f = open(self.path, "a", buffering=1 << 20)
        self.level = level

    def _emit(self, event: str, **kv):
        ts = datetime.datetime.now(datetime.timezone.utc).strftime("%Y-%m-%dT%H:%M:%S.%fZ")
        body = ",".join(f"{k}={str(v).replace(',', ';').replace(chr(10), ' ')}" for k, v in kv.items() if v is not None)
        self._f.write(f"{ts},{event},{body}\n")

    def job_start(self, plan):
        self._emit("start", job=os.path.basename(self.job_dir), stages=len(plan.nodes))
        for n in plan.nodes:
            self._emit("topology", stage=n.uid, stagename=n.name, type=n.type.value, partitions=n.partition_count,
                       children=" ".join(str(c.uid) for c, _ in n.children))
        with open(os.path.join(self.job_dir, "DryadLinqProgram__0.xml"), "w") as f:
            f.write(plan.to_xml())
        with open(os.path.join(self.job_dir, "QueryGraph__0.txt"), "w") as f:
            f.write(plan.explain())
        # closure objects the plan refers to by index travel with the job (DryadLinqObjectStore role)
        if hasattr(plan, "save_object_store"):
            plan.save_object_store(os.path.join(self.job_dir, "DryadLinqObjectStore__0.bin"))

    def vertex_event(self, node, part, version, state, **kv):
        self._emit("vertex", vertex=f"{node.uid}.{part}", version=version, stagename=node.name, stage=node.uid,
                   partInStage=part, state=state, **kv)

    def job_stop(self, ok: bool, error: str = "", **kv):
        self._emit("stop", success=ok, error=error or None, **kv)
        self._f.flush()
        if not ok and error:
            with open(os.path.join(self.job_dir, "log", "error.txt"), "w") as f:
                f.write(error)

    def close(self):
        self._f.close()
