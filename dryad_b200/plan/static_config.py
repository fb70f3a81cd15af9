"""Planner tunables (reference: StaticConfig, LinqToDryad/DryadLinqGlobals.cs:36-179)."""

# DynamicOptLevel bit flags (DryadLinqGlobals.cs:40-52); default = Broadcast only.
DYN_BROADCAST = 1
DYN_AGGREGATE = 2
DYN_RANGE_PARTITION = 4
DYN_HASH_PARTITION = 8
DYN_CO_RANGE = 16
DYNAMIC_OPT_LEVEL = DYN_BROADCAST

DEFAULT_PARTITION_COUNT = 8          # DryadLinqGlobals.cs:61
MAX_PARTITION_COUNT = 20000          # DryadLinqGlobals.cs:69
USE_MEMORY_FIFO = False              # DryadLinqGlobals.cs:74
USE_LARGE_BUFFER = False             # vertex write buffers: 256 MB instead of 8 MB (VertexEnv.write_buffer_size)
GROUP_BY_DYNAMIC_REDUCE = False      # tag partial GroupBy stages for the runtime aggregation tree even when
                                     # DYN_AGGREGATE is off (DryadLinqGlobals.cs:103)
GROUP_BY_LOCAL_AGGREGATION_IS_PARTIAL = True   # False: decomposable GroupBy shuffles raw rows and aggregates once
USE_PARTIAL_DISTINCT = False
SAMPLE_RATE = 0.001                  # DryadLinqSampler.cs:38
TEMP_LEASE_SECONDS = 24 * 3600       # DryadLinqGlobals.cs:143 (1 day, renewed every 23 h)
