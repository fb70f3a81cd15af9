"""CPU oracle: LocalDebug semantics of every operator (the correctness reference and the timed
CPU baseline).  Written from SURVEY.md; cites the reference behaviour it restates."""
